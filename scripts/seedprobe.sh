for g in 32 64 128; do
 CM_GPU_L2_FETCH=$g python bench.py --steps 5 --warmup 3 --upto 1 --no-cpu-baseline > gpurun_out/seed_g$g.json 2>gpurun_out/seed_g$g.err
 CM_GPU_L2_FETCH=$g ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sectors_srcunit_tex_op_read.sum,lts__t_sector_hit_rate.pct --clock-control none -k regex:^k_seed$ -c 1 --csv --log-file gpurun_out/seed_g$g.csv python bench.py --steps 1 --warmup 0 --upto 1 --no-cpu-baseline > /dev/null 2>&1
done
