# two GPUs: scene binaries under rtrun -n 2 (Main and DistributedMain), the full GPU suite, bench at N = 2
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
( time python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus 2 > gpurun_out/r2m_bench_n2.log 2> gpurun_out/r2m_bench_n2.err ) 2>&1 | grep real
tail -c 300 gpurun_out/r2m_bench_n2.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2m_bench_n2.log').read().strip().splitlines()[-1])
print('cornell n=2', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['ms_per_step'],2))
for k,v in d.get('workloads',{}).items(): print(k, round(v['value'],1), 'e2e', round(v['e2e']['value'],1), 'ms', round(v['ms_per_step'],2))
P
