for mb in 5 8 10 12; do
RT_WF_MINB=$mb timeout 900 python tools/perf_sweep.py --strategy wavefront --trav 0 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/WF_MINB=$mb /"
done
RT_MINB=8 timeout 900 python tools/perf_sweep.py --trav 0 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/fused MINB=8 /"
