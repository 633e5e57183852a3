# ncu captures of the path kernel on every workload class (one launch each, reduced spp)
cap() { W=$1; SPP=$2; timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_render -c 1 -f -o gpurun_out/r2n_$W \
    python bench.py --workload $W --spp $SPP --no-cpu-baseline --steps 1 --warmup 0 > gpurun_out/r2n_ncu_$W.log 2>&1; tail -1 gpurun_out/r2n_ncu_$W.log | cut -c1-80; }
cap birthday 32
cap bunny 64
cap spheres_d10 16
cap soup1m 8
cap soup10m 2
# launch list of the default bench command (share of the step per kernel)
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2n_launches.csv python bench.py --only-headline --no-cpu-baseline --steps 2 --warmup 1 > gpurun_out/r2n_launches.log 2>&1
tail -3 gpurun_out/r2n_launches.csv | cut -c1-200
