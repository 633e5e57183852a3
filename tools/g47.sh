timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
python __graft_entry__.py smoke 2>&1 | tail -1
python bench.py --workload bunny > gpurun_out/r1ae_bench_bunny.log 2>&1; tail -1 gpurun_out/r1ae_bench_bunny.log | cut -c1-100
python bench.py > gpurun_out/r1ae_bench_cornell.log 2>&1; tail -1 gpurun_out/r1ae_bench_cornell.log | cut -c1-100
C="python bench.py --workload bunny --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r1ae_bunny_launches.csv $C > gpurun_out/r1ae_ncu_l.log 2>&1; tail -1 gpurun_out/r1ae_ncu_l.log | cut -c1-100
