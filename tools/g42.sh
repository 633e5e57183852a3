python bench.py > gpurun_out/r1x_bench_cornell.log 2>&1; tail -1 gpurun_out/r1x_bench_cornell.log | cut -c1-200
python bench.py --workload birthday > gpurun_out/r1x_bench_birthday.log 2>&1; tail -1 gpurun_out/r1x_bench_birthday.log | cut -c1-200
C="python bench.py --steps 2 --warmup 3 --spp 32 --no-cpu-baseline"
$C > gpurun_out/r1x_plain.log 2>&1 && timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1x_cornell_launches.csv $C > gpurun_out/r1x_ncu_l.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1x_prof_cornell $C > gpurun_out/r1x_ncu_f.log 2>&1
tail -1 gpurun_out/r1x_ncu_f.log
