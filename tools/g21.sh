timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variants or degenerate" 2>&1 | tail -3
for mb in 5 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 0,1,10,15 --scenes cornell_box,spheres,bunny,birthday,soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
