set -x
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "variants" > gpurun_out/r1g7_pytest.log 2>&1; tail -5 gpurun_out/r1g7_pytest.log
timeout 900 python tools/perf_sweep.py --trav 1,12,13 --env "RT_MINB=4|5|6|8" --scenes cornell_box,spheres,birthday > gpurun_out/r1g7_sweep.txt 2>&1; grep -v "^{" gpurun_out/r1g7_sweep.txt | tail -40
timeout 900 python tools/perf_sweep.py --trav 0,12,13 --env "RT_MINB=4|5|8" --scenes bunny,soup1m > gpurun_out/r1g7_sweep2.txt 2>&1; grep -v "^{" gpurun_out/r1g7_sweep2.txt | tail -40
