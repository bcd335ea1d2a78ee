# A/B of launch-bounds / block-size variants of the library (built by hand next to libcm_gpu.so)
for v in "$@"; do
 CM_GPU_LIB=$PWD/circminer_b200/csrc/libcm_gpu_$v.so python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/lb_$v.json 2>gpurun_out/lb_$v.err
done
