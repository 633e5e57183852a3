tools/gpu_call.sh r2z tests bench w:spheres_d10:RT_BUILD_SMALL=0 w:spheres_d10 w:spheres_d50 w:cornell_box:RT_BUILD_SMALL=0 py:tools/launch_overhead.py
timeout 600 python tools/perf_sweep.py --trav 0,1 --scenes bunny1k,bunny3k,bunny6k,bunny20k,soup2k,soup6k,soup20k,soup100k,cornell54 2>&1 | grep -v "^{" | tee gpurun_out/r2z_inbetween_sweep.txt
