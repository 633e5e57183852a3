for d in 10 50; do
  python bench.py --workload spheres --depth $d --no-cpu-baseline > gpurun_out/r1ab_spheres_d${d}_auto.log 2>&1; echo "auto $(tail -1 gpurun_out/r1ab_spheres_d${d}_auto.log | cut -c1-90)"
  RT_SPLIT_MINCHUNK=1 RT_SPLIT=16 python bench.py --workload spheres --depth $d --no-cpu-baseline > gpurun_out/r1ab_spheres_d${d}_s16.log 2>&1; echo "S=16 $(tail -1 gpurun_out/r1ab_spheres_d${d}_s16.log | cut -c1-90)"
  RT_SPLIT_MINCHUNK=1 RT_SPLIT=16 RT_MINB=5 python bench.py --workload spheres --depth $d --no-cpu-baseline > gpurun_out/r1ab_spheres_d${d}_s16m5.log 2>&1; echo "S=16 minb5 $(tail -1 gpurun_out/r1ab_spheres_d${d}_s16m5.log | cut -c1-90)"
done
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
