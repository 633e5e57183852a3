set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r1g6_pytest.log 2>&1; tail -15 gpurun_out/r1g6_pytest.log
timeout 900 python tools/perf_sweep.py --trav 0,1,9,10,11 --scenes cornell_box,spheres,bunny,birthday,soup1m > gpurun_out/r1g6_sweep.txt 2>&1; grep -v "^{" gpurun_out/r1g6_sweep.txt | tail -30
