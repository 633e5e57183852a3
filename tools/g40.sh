timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "box_list or kernel_variants or render_bit_exact" 2>&1 | tail -3
for mb in 5 6; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 20 --scenes cornell_box,birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
