# round 2, first GPU call: parity of the additions made after the round-1 GPU budget was spent, then the
# cost-ordered tile schedule (RT_RENDER_ORDERED) against raster order on the bunny frame, with S = 1 / 4 / 32 slices.
timeout 1500 python -m pytest tests -m gpu -x -q 2>&1 | tail -4
for S in 1 4 32; do
  for O in "" "--ordered"; do
    RT_SPLIT=$S python bench.py --workload bunny --no-cpu-baseline $O > gpurun_out/r2a_bunny_S${S}${O:+_ordered}.log 2>&1
    tail -1 gpurun_out/r2a_bunny_S${S}${O:+_ordered}.log | cut -c1-120
  done
done
