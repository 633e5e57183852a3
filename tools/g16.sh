# image parity against the reference's CUDA render at 4096 spp with the loop-free sampler (north_star: rRMSE <= 1 %)
timeout 1500 python baseline/run_ref_cuda.py --tag r2 --scenes spheres,cornell_box,birthday --image-spp 4096 --timeout 500 2>&1 | cut -c1-1500
timeout 900 python baseline/run_ref_cuda.py --tag r2_bunny --scenes bunny --image-spp 256 --timeout 500 2>&1 | cut -c1-1500
# e2e anomaly check: birthday alone, with baselines
python bench.py --workload birthday --steps 3 > gpurun_out/r2s_birthday.log 2>&1; tail -1 gpurun_out/r2s_birthday.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('birthday', d['value'], d['e2e'])"
