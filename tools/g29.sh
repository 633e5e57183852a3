timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q 2>&1 | tail -3
timeout 900 python tools/perf_sweep.py --trav 0,1,17,18 --scenes cornell_box,spheres,bunny,birthday --reps 3 2>&1 | grep -v "^{"
