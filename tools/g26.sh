# compact wide tree (RT_TRAV 21/22): parity, then sweep against the current defaults
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "traversal_variants or degenerate_directions or compact_wide or render_kernel_variants" 2>&1 | tail -8
for mb in 8 6; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 0,16,21,22 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
RT_MINB=5 timeout 600 python tools/perf_sweep.py --trav 0,17,21,22 --scenes bunny,soup100k --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=5 /"
