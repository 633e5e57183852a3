timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/perf_sweep.py --trav 0,17 --scenes bunny,soup100k,soup10m8 --reps 3 2>&1 | grep -v "^{"
