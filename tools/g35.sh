for w in 4 1; do
RT_WPB=$w timeout 900 python tools/perf_sweep.py --trav 18 --scenes cornell_box --reps 3 2>&1 | grep -v "^{" | sed "s/^/WPB=$w /"
RT_WPB=$w timeout 900 python tools/perf_sweep.py --trav 1 --scenes spheres,birthday --reps 3 2>&1 | grep -v "^{" | sed "s/^/WPB=$w /"
RT_WPB=$w timeout 900 python tools/perf_sweep.py --trav 17 --scenes bunny --reps 3 2>&1 | grep -v "^{" | sed "s/^/WPB=$w /"
RT_WPB=$w timeout 900 python tools/perf_sweep.py --trav 0 --scenes soup1m16,soup10m8 --reps 2 2>&1 | grep -v "^{" | sed "s/^/WPB=$w /"
done
