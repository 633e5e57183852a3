nvidia-smi -L | wc -l
run() { n=$1; w=$2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 295$3 bench.py --gpus $n --steps 3 --warmup 3 --workload $w > gpurun_out/r1m_scale_${w}_$n.log 2>&1; tail -1 gpurun_out/r1m_scale_${w}_$n.log | cut -c1-260; }
run 8 cornell_box 11
run 8 birthday 12
run 4 birthday 13
run 2 birthday 14
run 8 soup10m 15
