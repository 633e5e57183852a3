# streaming kernel with batched service of the lanes between two queries (threshold sweep)
RT_TRAV=0 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "render_bit_exact or work_distribution" 2>&1 | tail -2
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2q_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2q_${W}_$tag.log | cut -c1-70)"; }
for W in soup1m soup10m; do for t in 4 8 16 24; do b thr$t env RT_STREAM_THR=$t; done; b query_loop env RT_STREAM=0; done
