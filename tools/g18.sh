# PLOC upper tree against the LBVH on the tree scenes
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_device_reference.py -m gpu -q 2>&1 | tail -3
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2u_${W}_$tag.log 2>&1; python - gpurun_out/r2u_${W}_$tag.log $W $tag <<'P'
import json,sys
d=json.loads(open(sys.argv[1]).read().strip().splitlines()[-1])
print(sys.argv[2], sys.argv[3], round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'box/ray', round(d['roofline']['algorithmic']['box_tests_per_ray'],1), 'rec/ray', round(d['roofline']['algorithmic']['record_tests_per_ray'],2), d['build'], 'nodes', d['config']['nodes'])
P
}
for W in soup1m soup10m bunny; do b ploc env; b lbvh env RT_PLOC=0; done
