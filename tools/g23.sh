# scaling of the headline (and birthday) at 8 and 4 GPUs, both reduces
for n in 8 4; do
for r in peers nccl; do
  timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2955$n \
    bench.py --gpus $n --only-headline --no-cpu-baseline --reduce $r 2> gpurun_out/g23_${n}_$r.err | tail -1 > gpurun_out/g23_bench_n${n}_$r.json
  python -c "
import json; d=json.loads(open('gpurun_out/g23_bench_n${n}_$r.json').read()); print($n, '$r', round(d['value'],1), 'ms', round(d['ms_per_step'],3), 'kernel', round(d['roofline']['kernel_ms'],3), 'e2e', round(d['e2e']['value'],1), round(d['e2e']['ms_per_step'],3), d['config'].get('image_crc32'))"
done
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29561 \
    bench.py --gpus 8 --workload birthday --no-cpu-baseline 2> gpurun_out/g23_birthday.err | tail -1 > gpurun_out/g23_bench_n8_birthday.json
python -c "
import json; d=json.loads(open('gpurun_out/g23_bench_n8_birthday.json').read()); print('birthday', d['value'], d['ms_per_step'], d['roofline']['kernel_ms'], d['e2e']['value'], d['config'].get('image_crc32'))"
