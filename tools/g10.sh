set -x
timeout 1200 python -m pytest tests -m gpu -x -q > gpurun_out/r1g10_pytest.log 2>&1; tail -8 gpurun_out/r1g10_pytest.log
