# peer-frame reduce on two GPUs: tests, the scene binary under rtrun, the headline at N=2 with both reduces
timeout 900 python -m pytest tests/test_peer_reduce.py tests/test_scene_api.py -m gpu -q 2>&1 | tail -6
for r in peers nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 \
    bench.py --gpus 2 --only-headline --no-cpu-baseline --reduce $r 2> gpurun_out/g22_$r.err | tail -1 > gpurun_out/g22_bench_n2_$r.json
  python -c "
import json; d=json.loads(open('gpurun_out/g22_bench_n2_$r.json').read()); print('$r', d['value'], d['ms_per_step'], d['e2e']['value'], d['e2e']['ms_per_step'], d['config'].get('image_crc32'), d['config'].get('reduce'))"
done
for w in birthday bunny; do
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 \
    bench.py --gpus 2 --workload $w --no-cpu-baseline 2> gpurun_out/g22_$w.err | tail -1 > gpurun_out/g22_bench_n2_$w.json
  python -c "
import json; d=json.loads(open('gpurun_out/g22_bench_n2_$w.json').read()); print('$w', d['value'], d['ms_per_step'], d['e2e']['value'], d['config'].get('image_crc32'))"
done
