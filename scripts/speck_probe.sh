# A/B: pre-alignment depth x heavy-pair warp kernel
for k in 1 2 4 8; do for t in 4 16; do
 CM_GPU_SPEC_K=$k CM_GPU_HEAVY_PAIR_TH=$t python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/speck_${k}_$t.json 2>/dev/null
done; done
