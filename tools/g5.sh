# round-1 session-20 re-baseline: GPU tests, headline bench, fresh ncu source capture of the default kernel
set -x
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r1g5_pytest.log 2>&1; tail -3 gpurun_out/r1g5_pytest.log
timeout 600 python bench.py > gpurun_out/r1g5_bench_cornell.log 2>&1; tail -1 gpurun_out/r1g5_bench_cornell.log | cut -c1-600
C="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
$C > gpurun_out/r1g5_plain.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1g5_prof_cornell $C > gpurun_out/r1g5_ncu.log 2>&1
ls -la gpurun_out | tail -5
