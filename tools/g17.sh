# breadth-first node layout + persisting L2 window on the large scenes
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q 2>&1 | tail -3
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2t_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2t_${W}_$tag.log | cut -c1-70)"; }
for W in soup1m soup10m; do
  b bfs_w48 env
  b bfs_nowin env RT_L2_WINDOW_MB=0
  b bfs_w24 env RT_L2_WINDOW_MB=24
  b bfs_w80 env RT_L2_WINDOW_MB=80
  b karras env RT_BFS=0
done
W=bunny; b bfs env; b karras env RT_BFS=0
