timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "render_bit_exact or full_size or smoke" 2>&1 | tail -3
for s in wavefront wavefront_nosort; do
timeout 900 python tools/perf_sweep.py --strategy $s --trav 1 --scenes cornell_box,spheres,bunny,birthday --reps 2 2>&1 | grep -v "^{" | sed "s/^/$s /"
done
