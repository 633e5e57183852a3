"""Fixed cost of one k_render launch (diagnostic, one GPU): Cornell 1024x1024 at spp = 8 .. 1024, kernel time from the
events inside rt_render_fixed (memset of the sums + kernel), L2 flushed between launches; least-squares fit
t = c + k * spp per work plan.  The 8-GPU headline renders 128 spp per GPU: c is what keeps it below 8.0x.
usage: python tools/launch_overhead.py <tag>"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
rt = importlib.import_module("ray-tracing-cuda_b200")

tag = sys.argv[1] if len(sys.argv) > 1 else "x"
ctx = rt.Context(0)
scene, cam = rt.scenes.cornell_box()
ctx.commit(scene)
ctx.set_camera(rt.core.camera_look_at(cam))
H = W = 1024
sums = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
plans = {"default": {}, "chunk16x1": {"RT_CHUNK": "16", "RT_ITEM_CHUNKS": "1"}, "chunk4x1": {"RT_CHUNK": "4", "RT_ITEM_CHUNKS": "1"},
         "chunk16x4": {"RT_CHUNK": "16", "RT_ITEM_CHUNKS": "4"}, "noflush": {}}
spps = [8, 16, 32, 64, 128, 256, 512, 1024]
out = {}
for name, env in plans.items():
    for k in ("RT_CHUNK", "RT_ITEM_CHUNKS"):
        os.environ.pop(k, None)
    os.environ.update(env)
    ts = []
    for spp in spps:
        best = []
        for rep in range(4):
            if name != "noflush":
                flush.fill_(1)
            torch.cuda.synchronize()
            st = ctx.render_fixed_device(sums.data_ptr(), H, W, spp, 0, 10, seed=1, flags=rt.abi.RT_RENDER_CLEAR)
            best.append(st.render_ms)
        ts.append(sorted(best[1:])[1])  # median of the last three
    A = np.stack([np.ones(len(spps)), np.array(spps, float)], 1)
    (c, k), *_ = np.linalg.lstsq(A, np.array(ts), rcond=None)
    (c2, k2), *_ = np.linalg.lstsq(A[3:], np.array(ts[3:]), rcond=None)
    out[name] = {"ms": dict(zip(map(str, spps), ts)), "fit_all": [c, k], "fit_64_up": [c2, k2]}
    print(f"{name:10s} " + " ".join(f"{s}:{t:.3f}" for s, t in zip(spps, ts)) + f"   c={c:.3f} ms k={k:.5f} ms/spp   (spp>=64: c={c2:.3f} k={k2:.5f})", flush=True)
os.makedirs(os.path.join(ROOT, "gpurun_out"), exist_ok=True)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"{tag}_launch_overhead.json"), "w"))
