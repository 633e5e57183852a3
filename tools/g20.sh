# early split clipping of triangles: soups + bunny at RT_TRI_SPLIT = 1, 2, 4; traversal parity with splitting on
for k in 1 2 4; do
  for w in soup1m bunny soup10m; do
    echo "== split $k $w"
    RT_TRI_SPLIT=$k timeout 600 python bench.py --workload $w --only-headline 2> /dev/null | python -c "
import sys, json
d = json.loads(sys.stdin.read().strip().splitlines()[-1])
print(round(d['value'],1), 'e2e', round(d['e2e']['value'],1), 'ms', round(d['ms_per_step'],2), d.get('build'), d['roofline']['algorithmic'])
"
  done
done
RT_TRI_SPLIT=3 timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x 2>&1 | tail -3
