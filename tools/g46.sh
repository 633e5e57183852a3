B="python bench.py --no-cpu-baseline --steps 3"
for s in 8 16 32; do RT_SPLIT=$s $B --workload bunny > gpurun_out/r1ad_bunny_s$s.log 2>&1; echo "bunny S=$s $(tail -1 gpurun_out/r1ad_bunny_s$s.log | cut -c1-90)"; done
$B --workload bunny > gpurun_out/r1ad_bunny_auto.log 2>&1; echo "bunny auto $(tail -1 gpurun_out/r1ad_bunny_auto.log | cut -c1-90)"
for s in 16 32; do RT_SPLIT=$s $B > gpurun_out/r1ad_cornell_s$s.log 2>&1; echo "cornell S=$s $(tail -1 gpurun_out/r1ad_cornell_s$s.log | cut -c1-90)"; done
B="python bench.py --no-cpu-baseline --steps 2"
for s in 1 4 16; do RT_SPLIT=$s $B --workload soup1m > gpurun_out/r1ad_soup1m_s$s.log 2>&1; echo "soup1m S=$s $(tail -1 gpurun_out/r1ad_soup1m_s$s.log | cut -c1-90)"; done
for s in 1 8; do RT_SPLIT=$s $B --workload soup10m > gpurun_out/r1ad_soup10m_s$s.log 2>&1; echo "soup10m S=$s $(tail -1 gpurun_out/r1ad_soup10m_s$s.log | cut -c1-90)"; done
