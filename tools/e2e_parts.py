"""Where the end-to-end time of a small frame goes (diagnostic, one GPU): wall-clock of the steps of bench.py's e2e
step — rt_scene_commit, rt_render_fixed, rt_resolve_fixed + D2H — each followed by a synchronize, median of 20.
usage: python tools/e2e_parts.py <tag>"""
import importlib, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
rt = importlib.import_module("ray-tracing-cuda_b200")
ctx = rt.Context(0)
CASES = {"spheres": (225, 400, 16), "cornell_box": (256, 256, 16), "bunny": (1080, 1920, 8)}
for name, (H, W, spp) in CASES.items():
    scene, cam = rt.scenes.by_name(name)
    camera = rt.core.camera_look_at(cam)
    sums = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    rgb = torch.zeros(H, W, 3, device="cuda")
    host = torch.zeros(H, W, 3).pin_memory()
    sp = torch.cuda.current_stream().cuda_stream
    t = {"commit": [], "commit_build_ms": [], "render": [], "render_kernel_ms": [], "resolve_d2h": [], "all": []}
    for rep in range(25):
        torch.cuda.synchronize()
        t0 = time.perf_counter()
        bs = ctx.commit(scene)
        ctx.set_camera(camera)
        torch.cuda.synchronize()
        t1 = time.perf_counter()
        st = ctx.render_fixed_device(sums.data_ptr(), H, W, spp, 0, 10, seed=1, flags=rt.abi.RT_RENDER_CLEAR, stream=sp)
        torch.cuda.synchronize()
        t2 = time.perf_counter()
        ctx.resolve_fixed_device(sums.data_ptr(), H, W, spp, None, rgb.data_ptr(), None, sp)
        host.copy_(rgb, non_blocking=True)
        torch.cuda.synchronize()
        t3 = time.perf_counter()
        if rep >= 5:
            t["commit"].append((t1 - t0) * 1e3); t["commit_build_ms"].append(bs.build_ms)
            t["render"].append((t2 - t1) * 1e3); t["render_kernel_ms"].append(st.render_ms)
            t["resolve_d2h"].append((t3 - t2) * 1e3); t["all"].append((t3 - t0) * 1e3)
    print(f"{name:12s} {W}x{H}x{spp}: " + "  ".join(f"{k} {np.median(v):.3f}" for k, v in t.items()), flush=True)
