timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_kernel_variants" 2>&1 | tail -2
for mb in 5 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 0,14 --scenes soup1m16,soup10m8,bunny --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
for th in "1 1" "4 4" "16 16" "4 16" "2 8"; do set -- $th
RT_MINB=8 RT_STREAM_LEAF=$1 RT_STREAM_SHADE=$2 timeout 900 python tools/perf_sweep.py --trav 14 --scenes soup1m16 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=8 L=$1 S=$2 /"
done
