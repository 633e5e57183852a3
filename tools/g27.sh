run() { n=$1; w=$2; python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 296$3 bench.py --gpus $n --steps 3 --warmup 3 --workload $w > gpurun_out/r1r_scale_${w}_$n.log 2>&1; tail -1 gpurun_out/r1r_scale_${w}_$n.log | cut -c1-200; }
run 8 cornell_box 11
run 8 bunny 12
run 4 cornell_box 13
run 2 cornell_box 14
