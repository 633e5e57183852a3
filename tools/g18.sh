for w in bunny birthday; do timeout 900 python bench.py --workload $w --steps 3 --warmup 3 > gpurun_out/r1k_bench_$w.log 2>&1; tail -1 gpurun_out/r1k_bench_$w.log | cut -c1-250; done
timeout 600 python bench.py --impl reference --workload bunny --steps 2 --warmup 1 2>&1 | tail -1 | cut -c1-600
