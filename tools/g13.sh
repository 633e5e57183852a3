# streaming kernel (one traversal step per iteration) against the per-query loop on the soups
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_zz_analytic_kat.py -m gpu -q 2>&1 | tail -3
RT_TRAV=0 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "render_bit_exact or work_distribution or textured or depth" 2>&1 | tail -3
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2p_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2p_${W}_$tag.log | cut -c1-70)"; }
for W in soup1m soup10m; do
  b stream env
  b stream_minb6 env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_st6.so
  b query_loop env RT_STREAM=0
done
W=bunny; b trav0_stream env RT_TRAV=0; b trav0_query env RT_TRAV=0 RT_STREAM=0; b default env
