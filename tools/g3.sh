# second cut: 4-sample chunks / one-chunk items, self-box-aware candidate order; the full default bench line
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
( time python bench.py > gpurun_out/r2e_bench_default.log 2> gpurun_out/r2e_bench_default.err ) 2>&1 | grep real
tail -c 600 gpurun_out/r2e_bench_default.err
python - <<'P'
import json
d=json.loads(open('gpurun_out/r2e_bench_default.log').read().strip().splitlines()[-1])
print('cornell', round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'frac', round(d['roofline']['frac'],3), d.get('ref_cuda_baseline'))
for k,v in d.get('workloads',{}).items():
    print(k, round(v['value'],1), 'e2e', round(v['e2e']['value'],1), 'frac', round(v['roofline']['frac'],3), 'ms', round(v['ms_per_step'],2), v.get('ref_cuda_baseline',{}).get('value'))
print(json.dumps(d.get('hit_bit_exact'), indent=None))
P
for v in u1 u4; do
  RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_$v.so python bench.py --workload cornell_box --no-cpu-baseline --steps 3 > gpurun_out/r2e_cornell_$v.log 2>&1; tail -1 gpurun_out/r2e_cornell_$v.log | cut -c1-120
done
