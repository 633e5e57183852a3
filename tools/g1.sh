# first run of the persistent path kernel (dynamic work items, fixed-point sums, loop-free sampler)
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -15
python __graft_entry__.py smoke 2>&1 | tail -2
for w in cornell_box birthday bunny spheres; do
  python bench.py --workload $w --no-cpu-baseline --steps 3 > gpurun_out/r2c_$w.log 2>&1; tail -1 gpurun_out/r2c_$w.log | cut -c1-200
done
python bench.py --workload spheres --depth 50 --no-cpu-baseline --steps 3 > gpurun_out/r2c_spheres50.log 2>&1; tail -1 gpurun_out/r2c_spheres50.log | cut -c1-200
for w in soup1m soup10m; do
  python bench.py --workload $w --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2c_$w.log 2>&1; tail -1 gpurun_out/r2c_$w.log | cut -c1-200
done
