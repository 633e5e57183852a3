# seventh cut: tree scenes single-phase; wavefront re-judged on soups (batched driver, optional ray reordering)
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
b() { tag=$1; shift; python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 "$@" > gpurun_out/r2l_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2l_${W}_$tag.log | cut -c1-70)"; }
for W in soup1m soup10m; do b fused; b wavefront --strategy wavefront; b wavefront_raysort --strategy wavefront_raysort; done
W=bunny; b fused; b wavefront --strategy wavefront
W=cornell_box; b fused; b wavefront --strategy wavefront
timeout 600 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o gpurun_out/r2l_cornell \
    python bench.py --workload cornell_box --spp 32 --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/r2l_ncu_cornell.log 2>&1
