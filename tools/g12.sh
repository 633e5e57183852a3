# shared-memory traversal stack on the soups: 16 levels (default build), 0 (all local), 24
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2o_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2o_${W}_$tag.log | cut -c1-70)"; }
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -k "soup or full_size or wide_tree or variants" 2>&1 | tail -3
for W in soup1m soup10m; do
  b sm16 env
  b sm0 env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_sm0.so
  b sm24 env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_sm24.so
done
