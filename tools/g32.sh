timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render" 2>&1 | tail -2
RT_MINB=8 timeout 900 python tools/perf_sweep.py --trav 0,16,17 --scenes soup1m16,soup10m8,bunny --reps 3 2>&1 | grep -v "^{" | sed "s/^/ool MINB=8 /"
