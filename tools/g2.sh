set -x
C="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
$C > gpurun_out/r1_plain_cornell.log 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r1_launches_cornell.csv $C > gpurun_out/r1_ncu_l_cornell.log 2>&1
ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1_prof_cornell $C > gpurun_out/r1_ncu_f_cornell.log 2>&1
B="python bench.py --workload bunny --spp 8 --steps 2 --warmup 3 --no-cpu-baseline"
$B > gpurun_out/r1_plain_bunny.log 2>&1 && ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1_prof_bunny $B > gpurun_out/r1_ncu_f_bunny.log 2>&1
for w in spheres bunny birthday; do for s in fused wavefront; do
  timeout 300 python bench.py --workload $w --strategy $s --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_bench_${w}_${s}.log 2>&1; tail -1 gpurun_out/r1_bench_${w}_${s}.log | cut -c1-400
done; done
timeout 300 python bench.py --strategy wavefront --spp 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r1_bench_cornell_wavefront.log 2>&1; tail -1 gpurun_out/r1_bench_cornell_wavefront.log | cut -c1-300
