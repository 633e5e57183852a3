timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "split" 2>&1 | tail -3
for s in 1 2 4 8; do RT_SPLIT=$s python bench.py --no-cpu-baseline --steps 3 > gpurun_out/r1ac_cornell_s$s.log 2>&1; echo "cornell S=$s $(tail -1 gpurun_out/r1ac_cornell_s$s.log | cut -c1-90)"; done
for s in 1 2 4; do RT_SPLIT=$s python bench.py --workload bunny --no-cpu-baseline --steps 3 > gpurun_out/r1ac_bunny_s$s.log 2>&1; echo "bunny S=$s $(tail -1 gpurun_out/r1ac_bunny_s$s.log | cut -c1-90)"; done
for s in 1 4; do RT_SPLIT=$s python bench.py --workload birthday --no-cpu-baseline --steps 3 > gpurun_out/r1ac_birthday_s$s.log 2>&1; echo "birthday S=$s $(tail -1 gpurun_out/r1ac_birthday_s$s.log | cut -c1-90)"; done
python bench.py --workload spheres --depth 50 > gpurun_out/r1ac_bench_spheres_d50.log 2>&1; tail -1 gpurun_out/r1ac_bench_spheres_d50.log | cut -c1-100
python bench.py --workload spheres > gpurun_out/r1ac_bench_spheres_d10.log 2>&1; tail -1 gpurun_out/r1ac_bench_spheres_d10.log | cut -c1-100
C="python bench.py --workload spheres --depth 50 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 300 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1ac_spheres_launches.csv $C > gpurun_out/r1ac_ncu_l.log 2>&1; tail -2 gpurun_out/r1ac_ncu_l.log | cut -c1-100
