# A/B of the warp-chaining threshold; also T1 (two contigs, live reference) with every strand on the warp kernel
CM_GPU_HEAVY_TH=0 python -m pytest tests/test_gpu_pipeline.py -x -q -k t1_live > gpurun_out/heavy_t1.log 2>&1
for t in 16 32 64 128 254; do
 CM_GPU_HEAVY_TH=$t python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/heavy_$t.json 2>gpurun_out/heavy_$t.err
done
