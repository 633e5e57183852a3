for mb in 6; do
RT_MINB=$mb RT_STREAM_LEAF=4 RT_STREAM_SHADE=4 timeout 900 python tools/perf_sweep.py --trav 0,14 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb L4S4 /"
RT_MINB=$mb RT_STREAM_LEAF=2 RT_STREAM_SHADE=8 timeout 900 python tools/perf_sweep.py --trav 14 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/MINB=$mb L2S8 /"
done
