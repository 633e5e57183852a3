# sixth cut: diagonal-edge parallelograms resolved without the sequential scan
timeout 1200 python -m pytest tests -m gpu -q 2>&1 | tail -4
b() { tag=$1; shift; "$@" python bench.py --workload $W --no-cpu-baseline --steps 3 > gpurun_out/r2j_${W}_$tag.log 2>&1; echo "$W $tag $(tail -1 gpurun_out/r2j_${W}_$tag.log | cut -c1-70)"; }
for W in bunny; do b default env; b ch4 env RT_CHUNK=4; b ch16r16 env RT_CHUNK=16 RT_ITEM_CHUNKS=16; done
for W in cornell_box birthday spheres_d10 soup1m soup10m; do b default env; done
bash tools/sanitize.sh > gpurun_out/r2j_sanitize.log 2>&1; grep -E "ERROR SUMMARY|exit code|====|sanitize case" gpurun_out/r2j_sanitize.log
