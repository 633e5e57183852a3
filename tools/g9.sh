set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/r1g9_pytest.log 2>&1; tail -5 gpurun_out/r1g9_pytest.log
timeout 900 python tools/perf_sweep.py --trav 0,1 --scenes cornell_box,spheres,bunny,birthday,soup1m > gpurun_out/r1g9_sweep.txt 2>&1; grep -v "^{" gpurun_out/r1g9_sweep.txt | tail -40
