timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -3
timeout 900 python bench.py --workload bunny --steps 3 --warmup 3 > gpurun_out/r1p_bench_bunny.log 2>&1; tail -1 gpurun_out/r1p_bench_bunny.log | cut -c1-250
timeout 900 python tools/perf_sweep.py --trav 0,1,17 --scenes soup100k --reps 2 2>&1 | grep -v "^{"
B="python bench.py --workload bunny --spp 8 --steps 2 --warmup 3 --no-cpu-baseline"
timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1p_prof_bunny $B > gpurun_out/r1p_ncu_b.log 2>&1
