for mb in 6 8; do
RT_MINB=$mb timeout 900 python tools/perf_sweep.py --trav 1,18 --scenes cornell_box,spheres,birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/noinline-sphere MINB=$mb /"
done
