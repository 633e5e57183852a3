"""Build time (rt_scene_commit, best of 3) and device memory held per scene, for the default traversal and with a wide
/ compact tree requested (RT_TRAV=17 / 22).  usage: python tools/build_times.py [scene ...]"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
rt = importlib.import_module("ray-tracing-cuda_b200")

def scene_of(name):
    if name.startswith("soup"):
        return rt.scenes.soup({"soup100k": 100000, "soup1m": 1000000, "soup10m": 10000000}[name])
    return rt.scenes.by_name(name)

for name in sys.argv[1:] or ["cornell_box", "spheres", "bunny", "soup1m", "soup10m"]:
    scene, cam = scene_of(name)
    for trav in (None, "17", "22"):
        if trav is None:
            os.environ.pop("RT_TRAV", None)
        else:
            os.environ["RT_TRAV"] = trav
        ctx = rt.Context(0)
        torch.cuda.synchronize()
        free0 = torch.cuda.mem_get_info()[0]
        best = min(ctx.commit(scene).build_ms for _ in range(3))
        bs = ctx.commit(scene)
        torch.cuda.synchronize()
        held = (free0 - torch.cuda.mem_get_info()[0]) / 2**20
        print(f"{name:12s} RT_TRAV={trav or 'default':8s} build {best:8.2f} ms  nodes {bs.n_nodes}  wide nodes {bs.n_wide_nodes}  "
              f"device memory held {held:8.1f} MiB", flush=True)
        ctx.close()
