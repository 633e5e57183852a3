timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_kernel_variants" 2>&1 | tail -2
for mb in 5 8; do
RT_MINB=$mb RT_STREAM_LEAF=2 RT_STREAM_SHADE=8 timeout 900 python tools/perf_sweep.py --trav 0,14 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb L2S8 /"
done
RT_MINB=8 RT_STREAM_LEAF=4 RT_STREAM_SHADE=4 timeout 900 python tools/perf_sweep.py --trav 14 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=8 L4S4 /"
