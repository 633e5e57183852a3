# fifth cut: render skips pow unless (float)t is undecided; tile order / item size experiments; footprint variants
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 3 > gpurun_out/r2h_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2h_${W}_$tag.log | cut -c1-70)"; }
for W in cornell_box birthday spheres_d10 spheres_d50; do b default env; done
for W in bunny soup1m; do
  b default env
  b golden env RT_TILE_ORDER=1
  b ch4 env RT_CHUNK=4
  b ch4_golden env RT_CHUNK=4 RT_TILE_ORDER=1
  b ch16r4_golden env RT_CHUNK=16 RT_ITEM_CHUNKS=4 RT_TILE_ORDER=1
done
W=cornell_box
b golden env RT_TILE_ORDER=1
b philox_loop env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_pl.so
b quad_loop env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_ql.so
W=birthday
b quad_loop env RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_ql.so
