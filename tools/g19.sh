b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2v_${W}_$tag.log 2>&1; python - gpurun_out/r2v_${W}_$tag.log $W $tag <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], sys.argv[3], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'box/ray', round(d['roofline']['algorithmic']['box_tests_per_ray'],1), 'rec/ray', round(d['roofline']['algorithmic']['record_tests_per_ray'],2), d['build'], 'nodes', d['config']['nodes'])
P
}
for W in soup1m bunny; do b r16 env RT_PLOC_R=16; b r32 env RT_PLOC_R=32; done
W=bunny; b r4 env RT_PLOC_R=4
W=bunny; b leaf7 env RT_MAX_LEAF=7
W=bunny; b leaf2 env RT_MAX_LEAF=2
