set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r1k_pytest.log 2>&1; tail -3 gpurun_out/r1k_pytest.log
timeout 900 python bench.py > gpurun_out/r1k_bench_cornell.log 2>&1; tail -1 gpurun_out/r1k_bench_cornell.log | cut -c1-300
timeout 600 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/r1k_bench_reference.log 2>&1; tail -1 gpurun_out/r1k_bench_reference.log | cut -c1-400
for w in spheres bunny birthday; do timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/r1k_bench_$w.log 2>&1; tail -1 gpurun_out/r1k_bench_$w.log | cut -c1-200; done
C="python bench.py --steps 2 --warmup 1 --no-cpu-baseline"
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1k_cornell_launches.csv $C > gpurun_out/r1k_ncu_l.log 2>&1
C="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1k_prof_cornell $C > gpurun_out/r1k_ncu_f.log 2>&1
B="python bench.py --workload bunny --spp 8 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1k_prof_bunny $B > gpurun_out/r1k_ncu_b.log 2>&1
ls gpurun_out | grep r1k
