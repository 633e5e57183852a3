timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_kernel_variants" 2>&1 | tail -3
for mb in 5 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 0,16,17 --scenes cornell_box,spheres,bunny,birthday,soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
