cp ray-tracing-cuda_b200/librtcore.so /tmp/librtcore_default.so
for v in default u3 u6; do
  if [ $v = default ]; then cp /tmp/librtcore_default.so ray-tracing-cuda_b200/librtcore.so; else cp gpuvar/librtcore_$v.so ray-tracing-cuda_b200/librtcore.so; fi
  timeout 600 python tools/perf_sweep.py --trav 20 --scenes cornell_box,birthday --reps 4 2>&1 | grep -v "^{" | sed "s/^/$v /"
done
cp /tmp/librtcore_default.so ray-tracing-cuda_b200/librtcore.so
RT_MINB=7 timeout 600 python tools/perf_sweep.py --trav 20 --scenes cornell_box --reps 4 2>&1 | grep -v "^{" | sed "s/^/MINB=7 /"
