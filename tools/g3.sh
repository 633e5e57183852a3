set -x
python tools/perf_sweep.py --trav 0,1 --scenes soup1m4k,soup10m --env "RT_MINB=5|8" --reps 2 2>&1 | grep -v "^{" | cut -c1-180 | tee gpurun_out/r1_sweep_soup.log
C="python tools/perf_sweep.py --trav 0 --scenes soup10m --reps 1"
$C > gpurun_out/r1_plain_soup10m.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 1 -c 1 -o gpurun_out/r1_prof_soup10m $C > gpurun_out/r1_ncu_soup10m.log 2>&1
C2="python bench.py --spp 32 --steps 1 --warmup 3 --no-cpu-baseline"
$C2 > gpurun_out/r1_plain_cornell2.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1_prof_cornell2 $C2 > gpurun_out/r1_ncu_cornell2.log 2>&1
ls -la gpurun_out | tail -5
