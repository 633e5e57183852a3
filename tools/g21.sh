# peer-frame reduce, one GPU: the world_size-1 tests + the headline line (image crc for the N-GPU comparison)
timeout 600 python -m pytest tests/test_peer_reduce.py -m gpu -q 2>&1 | tail -5
timeout 600 python bench.py --only-headline --no-cpu-baseline 2> gpurun_out/g21_bench.err | tail -1 > gpurun_out/g21_bench_n1.json
python -c "
import json; d=json.loads(open('gpurun_out/g21_bench_n1.json').read()); print(d['value'], d['e2e']['value'], d['config'].get('image_crc32'))"
