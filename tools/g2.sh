# ncu captures of the path kernel: bunny (regression vs sample slices?) and Cornell
for w in bunny cornell_box; do
  timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o gpurun_out/r2d_$w \
    python bench.py --workload $w --spp 32 --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/r2d_ncu_$w.log 2>&1
  tail -2 gpurun_out/r2d_ncu_$w.log | cut -c1-300
done
ls -la gpurun_out/*.ncu-rep
