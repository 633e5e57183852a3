# cooperative ball rejection (RT_TRAV 23) against the box list (20)
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "render_kernel_variants or render_bit_exact" 2>&1 | tail -5
for mb in 6 5; do
RT_MINB=$mb timeout 600 python tools/perf_sweep.py --trav 20,23,20,23 --scenes cornell_box,birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
