# third cut: chunk rule per_lane/256 (<= 16), box-list unroll 1, self-box order, kernels specialised on sphere presence,
# Sphere::Hit through CUDA's pow (reference-exact) with an IEEE pre-filter
timeout 1200 python -m pytest tests -m gpu -x -q 2>&1 | tail -5
for w in cornell_box birthday bunny spheres_d10 soup1m; do
  python bench.py --workload $w --no-cpu-baseline --steps 3 > gpurun_out/r2f_$w.log 2>&1; tail -1 gpurun_out/r2f_$w.log | cut -c1-110
done
RT_SPHERE_POW=ieee python bench.py --workload spheres_d10 --no-cpu-baseline --steps 3 > gpurun_out/r2f_spheres_ieee.log 2>&1; tail -1 gpurun_out/r2f_spheres_ieee.log | cut -c1-110
RT_SPHERE_POW=ieee python bench.py --workload birthday --no-cpu-baseline --steps 3 > gpurun_out/r2f_birthday_ieee.log 2>&1; tail -1 gpurun_out/r2f_birthday_ieee.log | cut -c1-110
for w in cornell_box birthday; do
  RT_CORE_LIB=$PWD/ray-tracing-cuda_b200/librtcore_noself.so python bench.py --workload $w --no-cpu-baseline --steps 3 > gpurun_out/r2f_${w}_noself.log 2>&1; tail -1 gpurun_out/r2f_${w}_noself.log | cut -c1-110
done
for c in 4 8 32; do
  RT_CHUNK=$c python bench.py --workload cornell_box --no-cpu-baseline --steps 3 > gpurun_out/r2f_cornell_ch$c.log 2>&1; tail -1 gpurun_out/r2f_cornell_ch$c.log | cut -c1-110
done
