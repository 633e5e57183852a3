timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --workload soup1m --steps 3 --warmup 3 > gpurun_out/r1s_bench_soup1m.log 2>&1; tail -1 gpurun_out/r1s_bench_soup1m.log | cut -c1-220
timeout 900 python bench.py --workload soup10m --steps 3 --warmup 3 > gpurun_out/r1s_bench_soup10m.log 2>&1; tail -1 gpurun_out/r1s_bench_soup10m.log | cut -c1-220
timeout 600 python bench.py --workload spheres --depth 50 --steps 3 --warmup 3 > gpurun_out/r1s_bench_spheres_d50.log 2>&1; tail -1 gpurun_out/r1s_bench_spheres_d50.log | cut -c1-220
timeout 600 python bench.py --workload birthday --steps 3 --warmup 3 > gpurun_out/r1s_bench_birthday.log 2>&1; tail -1 gpurun_out/r1s_bench_birthday.log | cut -c1-220
