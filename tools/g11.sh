set -x
timeout 1500 python baseline/run_ref_cuda.py --tag r1_4096 --image-spp 4096 --timeout 400 > gpurun_out/r1g11_ref.log 2>&1; tail -5 gpurun_out/r1g11_ref.log | cut -c1-900
