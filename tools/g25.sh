timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py > gpurun_out/r1q_bench_cornell.log 2>&1; tail -1 gpurun_out/r1q_bench_cornell.log | cut -c1-250
C="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1q_cornell_launches.csv $C > gpurun_out/r1q_ncu_l.log 2>&1
C="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1q_prof_cornell $C > gpurun_out/r1q_ncu_f.log 2>&1
