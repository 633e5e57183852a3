C="python bench.py --steps 2 --warmup 3 --spp 32"
$C > gpurun_out/r1w_plain_cornell.log 2>&1 && timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1w_prof_cornell $C > gpurun_out/r1w_ncu_f.log 2>&1
tail -2 gpurun_out/r1w_ncu_f.log
