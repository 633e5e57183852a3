# fourth cut: two-phase work plan, interval sphere test + one pow evaluation for the winner
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -15
for w in cornell_box birthday bunny spheres_d10 spheres_d50 soup1m soup10m; do
  python bench.py --workload $w --no-cpu-baseline --steps 3 > gpurun_out/r2g_$w.log 2>&1; tail -1 gpurun_out/r2g_$w.log | cut -c1-110
done
RT_SPHERE_POW=ieee python bench.py --workload spheres_d10 --no-cpu-baseline --steps 3 > gpurun_out/r2g_spheres_ieee.log 2>&1; tail -1 gpurun_out/r2g_spheres_ieee.log | cut -c1-110
RT_SPHERE_POW=ieee python bench.py --workload birthday --no-cpu-baseline --steps 3 > gpurun_out/r2g_birthday_ieee.log 2>&1; tail -1 gpurun_out/r2g_birthday_ieee.log | cut -c1-110
