timeout 900 python -m pytest tests/test_gpu_parity.py -x -q -m gpu -k "box_list or kernel_variants" 2>&1 | tail -5
for mb in 5 6 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 18,20 --scenes cornell_box --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 1,20 --scenes birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
