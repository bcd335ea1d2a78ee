# T1 (two contigs, live reference) with every pair on the warp kernel, then an A/B of the heavy-pair threshold
CM_GPU_HEAVY_PAIR_TH=0 python -m pytest tests/test_gpu_pipeline.py -x -q -k "t1_live or t2_k14" > gpurun_out/heavy_t1.log 2>&1
for t in 4 16 64 100000; do
 CM_GPU_HEAVY_PAIR_TH=$t python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/heavyp_$t.json 2>gpurun_out/heavyp_$t.err
done
