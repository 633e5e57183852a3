timeout 600 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "variants" 2>&1 | tail -3
RT_MINB=8 timeout 900 python tools/perf_sweep.py --trav 0,20,21,22 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{"
RT_MINB=5 timeout 900 python tools/perf_sweep.py --trav 0,20 --scenes bunny --reps 2 2>&1 | grep -v "^{"
