# after removing the wide tree and the streaming kernel: full GPU suite + the default bench line
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
( time python bench.py > gpurun_out/r2r_bench_default.log 2> gpurun_out/r2r_bench_default.err ) 2>&1 | grep real
tail -c 300 gpurun_out/r2r_bench_default.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2r_bench_default.log').read().strip().splitlines()[-1])
print('cornell', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), d.get('ref_cuda_baseline',{}).get('value'))
for k,v in d.get('workloads',{}).items():
    print(k, round(v['value'],1), 'e2e', round(v['e2e']['value'],1), 'frac', round(v['roofline']['frac'],3), 'ms', round(v['ms_per_step'],2), v.get('ref_cuda_baseline',{}).get('value'))
hb=d.get('hit_bit_exact',{})
for k,v in hb.items(): print(k, {a:b for a,b in v.items() if 'detail' not in a} if isinstance(v,dict) else v)
P
