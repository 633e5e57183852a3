for mb in 6 7; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 18 --scenes cornell_box --reps 4 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb /"
done
