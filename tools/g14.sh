set -x
C="python bench.py --workload soup1m --spp 8 --steps 1 --warmup 3 --no-cpu-baseline"
RT_MINB=5 RT_TRAV=14 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_stream -s 3 -c 1 -o gpurun_out/r1g14_prof_stream $C > gpurun_out/r1g14_ncu1.log 2>&1
RT_MINB=5 RT_TRAV=0 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render_fused -s 3 -c 1 -o gpurun_out/r1g14_prof_ifif $C > gpurun_out/r1g14_ncu2.log 2>&1
ls -la gpurun_out/*.ncu-rep | tail -3
