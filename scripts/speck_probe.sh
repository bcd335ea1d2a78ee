# A/B of the pre-alignment depth (CM_GPU_SPEC_K) on the bench workload
for k in 1 2 4 30; do
 CM_GPU_SPEC_K=$k python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/speck_$k.json 2>gpurun_out/speck_$k.err
done
