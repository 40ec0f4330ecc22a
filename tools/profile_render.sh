CMD="python bench.py --steps 1 --warmup 3 --no-cpu-baseline --no-inference --no-brute --no-extra-configs"
$CMD > gpurun_out/plain.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 300 --csv --log-file gpurun_out/r02_launches_bench_config3.csv $CMD > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:render_bvh_spp_kernel -s 4 -c 1 -f -o gpurun_out/r02_render_spp $CMD > gpurun_out/ncu_full.log 2>&1
tail -2 gpurun_out/ncu_full.log
