set -x
timeout 900 python bench.py --workload soup1m --steps 3 --warmup 3 > gpurun_out/r1g12_bench_soup1m.log 2>&1; tail -1 gpurun_out/r1g12_bench_soup1m.log | cut -c1-1500
timeout 900 python bench.py --workload soup10m --steps 3 --warmup 3 > gpurun_out/r1g12_bench_soup10m.log 2>&1; tail -1 gpurun_out/r1g12_bench_soup10m.log | cut -c1-1500
C="python bench.py --workload soup10m --spp 2 --steps 1 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1g12_prof_soup10m $C > gpurun_out/r1g12_ncu.log 2>&1; tail -2 gpurun_out/r1g12_ncu.log | cut -c1-300
