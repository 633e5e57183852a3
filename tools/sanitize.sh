#!/bin/bash
# compute-sanitizer over the small configuration (race detection / sanitizers): memcheck and racecheck of
# __graft_entry__.smoke() — commit + BVH build + closest hits + path kernel + wavefront pipeline + resolve on the
# Cornell box — plus one sphere scene and one mesh scene through the parity hook and both render strategies.
# Run on a GPU box:  bash tools/sanitize.sh > profiles/rNN_sanitize.log 2>&1
cd "$(dirname "$0")/.."
cat > /tmp/sanitize_case.py <<'P'
import importlib, sys, numpy as np
sys.path.insert(0, "."); sys.path.insert(0, "tests")
rt = importlib.import_module("ray-tracing-cuda_b200")
import __graft_entry__ as g
g.smoke()
ctx = rt.Context(0)
for name, hw in (("spheres", (27, 48)), ("bunny9k", (27, 48))):
    scene, cam = (rt.scenes.spheres(10086, "rtl")[:2] if name == "spheres" else rt.scenes.bunny(rt.scenes.bunny_mesh(9000)))
    ctx.commit(scene); ctx.set_camera(rt.core.camera_look_at(cam))
    rays = ctx.primary_rays(*hw)
    for mode in (0, 1, 2, 3):
        ctx.trace_closest(rays, mode=mode)
    for flags in (rt.abi.RT_RENDER_FUSED, rt.abi.RT_RENDER_WAVEFRONT, rt.abi.RT_RENDER_WAVEFRONT | rt.abi.RT_RENDER_SORT,
                  rt.abi.RT_RENDER_FUSED | rt.abi.RT_RENDER_PROFILE):
        ctx.render_host(hw[0], hw[1], 3, seed=5, flags=flags)
    print("sanitize case ok:", name)
ctx.close()
P
for tool in memcheck racecheck; do
  echo "==== compute-sanitizer --tool $tool"
  timeout 1500 compute-sanitizer --tool $tool --error-exitcode 9 python /tmp/sanitize_case.py 2>&1 | grep -v "^$" | tail -25
  echo "exit code: ${PIPESTATUS[0]}"
done
