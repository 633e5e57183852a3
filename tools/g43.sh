timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "split or bit_exact_vs_oracle or max_depth" 2>&1 | tail -4
for d in 10 50; do
  python bench.py --workload spheres --depth $d --no-cpu-baseline --no-split > gpurun_out/r1aa_spheres_d${d}_nosplit.log 2>&1; tail -1 gpurun_out/r1aa_spheres_d${d}_nosplit.log | cut -c1-120
  python bench.py --workload spheres --depth $d --no-cpu-baseline > gpurun_out/r1aa_spheres_d${d}_split.log 2>&1; tail -1 gpurun_out/r1aa_spheres_d${d}_split.log | cut -c1-120
  for s in 2 3 6 8; do RT_SPLIT=$s python bench.py --workload spheres --depth $d --no-cpu-baseline > gpurun_out/r1aa_spheres_d${d}_s$s.log 2>&1; echo "S=$s $(tail -1 gpurun_out/r1aa_spheres_d${d}_s$s.log | cut -c1-90)"; done
done
python bench.py --no-cpu-baseline --steps 3 > gpurun_out/r1aa_cornell.log 2>&1; tail -1 gpurun_out/r1aa_cornell.log | cut -c1-120
