# why does bunny prefer 4-sample chunks? ncu of both plans at the full 256 spp
for v in default ch4; do
  if [ $v = ch4 ]; then export RT_CHUNK=4; else unset RT_CHUNK; fi
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o gpurun_out/r2i_bunny_$v \
    python bench.py --workload bunny --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/r2i_ncu_$v.log 2>&1
  tail -1 gpurun_out/r2i_ncu_$v.log | cut -c1-100
done
