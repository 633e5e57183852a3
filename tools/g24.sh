timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_kernel_variants" 2>&1 | tail -3
for mb in 5 6 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 1,18,19 --scenes cornell_box,spheres,birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
