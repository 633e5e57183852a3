set -x
C="python bench.py --spp 32 --steps 2 --warmup 3 --no-cpu-baseline"
RT_TRAV=12 ncu --set full --clock-control none --import-source on -k regex:k_render_persist -s 3 -c 1 -o gpurun_out/r1g8_prof_persist $C > gpurun_out/r1g8_ncu.log 2>&1
RT_TRAV=10 ncu --set full --clock-control none --import-source on -k regex:k_render_warp -s 3 -c 1 -o gpurun_out/r1g8_prof_wide $C > gpurun_out/r1g8_ncu2.log 2>&1
ls -la gpurun_out | tail -4
