nvidia-smi -L | wc -l
for n in 8 4; do
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2951$n bench.py --gpus $n --steps 3 --warmup 3 > gpurun_out/r1_scale_cornell_$n.log 2>&1; tail -1 gpurun_out/r1_scale_cornell_$n.log | cut -c1-330
  python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2952$n bench.py --gpus $n --steps 3 --warmup 3 --workload bunny > gpurun_out/r1_scale_bunny_$n.log 2>&1; tail -1 gpurun_out/r1_scale_bunny_$n.log | cut -c1-330
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 --workload birthday > gpurun_out/r1_scale_birthday_8.log 2>&1; tail -1 gpurun_out/r1_scale_birthday_8.log | cut -c1-330
mkdir -p /tmp/rr && cd /tmp/rr && python $GRAFT_REPO_ROOT/tools/make_resources.py resources >/dev/null && RT_STATS=/tmp/rr/st.json RT_WIDTH=1920 RT_HEIGHT=1080 RT_SPP=256 $GRAFT_REPO_ROOT/ray-tracing-cuda_b200/scene_api/build/rtrun -n 8 $GRAFT_REPO_ROOT/ray-tracing-cuda_b200/scene_api/build/bunny 2>&1 | grep -i "finished\|workload" | head -16; cat /tmp/rr/st.json.0
