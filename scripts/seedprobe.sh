# A/B of the fused-directory load path (CM_GPU_DBG = 256 * variant): kernel time and DRAM bytes of k_seed
for g in 0 1 2 3 4 5; do
 CM_GPU_DBG=$((256*g)) python bench.py --steps 5 --warmup 3 --upto 1 --no-cpu-baseline > gpurun_out/seed_v$g.json 2>gpurun_out/seed_v$g.err
 CM_GPU_DBG=$((256*g)) ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sectors_srcunit_tex_op_read.sum --clock-control none -k regex:^k_seed$ -c 1 --csv --log-file gpurun_out/seed_v$g.csv python bench.py --steps 1 --warmup 0 --upto 1 --no-cpu-baseline > /dev/null 2>&1
done
