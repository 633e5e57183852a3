"""Kernel-variant sweep on one GPU (diagnostics, not the bench): Msamples/s of rt_render per scene and
RT_TRAV variant, plus node/record visits per ray from the stats build.
usage: python tools/perf_sweep.py [--trav 0,1,2,3,4] [--scenes cornell_box,spheres,bunny,birthday,soup1m]"""
import argparse, importlib, json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch
rt = importlib.import_module("ray-tracing-cuda_b200")

CASES = {"cornell_box": (1024, 1024, 64), "spheres": (720, 1280, 32), "bunny": (1080, 1920, 16),
         "birthday": (1080, 1920, 32), "soup1m": (1080, 1920, 4), "soup100k": (1080, 1920, 8), "soup10m": (2160, 3840, 2), "soup1m4k": (2160, 3840, 4), "soup1m16": (1080, 1920, 16), "soup10m8": (1080, 1920, 8),
         # scene classes between the named configs (the thresholds of default_trav): meshes and soups of 1 k - 100 k triangles,
         # a Cornell box of 54 parallelograms (three more blocks: no box list above 32 primitives)
         "bunny1k": (1080, 1920, 16), "bunny3k": (1080, 1920, 16), "bunny6k": (1080, 1920, 16), "bunny20k": (1080, 1920, 16),
         "soup2k": (1080, 1920, 8), "soup6k": (1080, 1920, 8), "soup20k": (1080, 1920, 8), "cornell54": (1024, 1024, 64)}

def scene_of(name):
    if name.startswith("soup"):
        n = {"soup2k": 2000, "soup6k": 6000, "soup20k": 20000, "soup1m": 1000000, "soup100k": 100000, "soup10m": 10000000, "soup1m4k": 1000000, "soup1m16": 1000000, "soup10m8": 10000000}[name]
        return rt.scenes.soup(n)
    if name.startswith("bunny") and name != "bunny":
        return rt.scenes.bunny(rt.scenes.bunny_mesh(int(name[5:-1]) * 1000))
    if name == "cornell54":
        s, cam = rt.scenes.cornell_box()
        rng = np.random.RandomState(4)
        hs = list(s.hitables)
        for b in range(6):  # six more small blocks on the floor, six parallelograms each
            o = np.array([rng.uniform(40, 480), 0.0, rng.uniform(40, 480)])
            e = [np.array([rng.uniform(30, 60), 0, rng.uniform(-20, 20)]), np.array([0, rng.uniform(30, 120), 0]), None]
            e[2] = np.array([-e[0][2], 0, e[0][0]])
            for a, b2, c in ((0, 1, 2), (1, 2, 0), (2, 0, 1)):
                for off in (np.zeros(3), e[c]):
                    hs.append(rt.scenes.parallelogram(np.stack([o + off, o + off + e[a], o + off + e[b2]]), 0))
        return rt.abi.SceneData(np.stack(hs), s.materials, name="cornell54"), cam
    return rt.scenes.by_name(name)

def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--trav", default="0,1,2,3,4")
    ap.add_argument("--scenes", default="cornell_box,spheres,bunny,birthday,soup1m")
    ap.add_argument("--strategy", default="fused")
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--env", default="", help="extra VAR=a|b|c axis")
    a = ap.parse_args()
    ctx = rt.Context(0)
    flags = {"fused": rt.abi.RT_RENDER_FUSED, "wavefront": rt.abi.RT_RENDER_WAVEFRONT,
             "wavefront_sorted": rt.abi.RT_RENDER_WAVEFRONT | rt.abi.RT_RENDER_SORT}[a.strategy]
    extra = [None]
    if a.env:
        var, vals = a.env.split("=")
        extra = [(var, v) for v in vals.split("|")]
    res = {}
    for name in a.scenes.split(","):
        H, W, spp = CASES[name]
        scene, cam = scene_of(name)
        ctx.set_camera(rt.core.camera_look_at(cam))
        acc = torch.zeros(H, W, 3, device="cuda")
        for ex in extra:
            if ex:
                os.environ[ex[0]] = ex[1]
            for tv in a.trav.split(","):
                os.environ["RT_TRAV"] = tv  # read at commit
                bs = ctx.commit(scene)
                st = ctx.render_device(acc.data_ptr(), H, W, 1, 0, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR | rt.abi.RT_RENDER_PROFILE)
                nodes, prims, rays1 = st.node_visits, st.prim_tests, st.rays
                simt = 0.0
                best = 1e30
                for _ in range(a.reps):
                    st = ctx.render_device(acc.data_ptr(), H, W, spp, 0, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR)
                    best = min(best, st.render_ms)
                key = f"{name} trav={tv}" + (f" {ex[0]}={ex[1]}" if ex else "")
                res[key] = dict(msamples=H * W * spp / best / 1e3, mrays=st.rays / best / 1e3, ms=best,
                                nodes_per_ray=nodes / max(rays1, 1), recs_per_ray=prims / max(rays1, 1),
                                build_ms=bs.build_ms, nodes=bs.n_nodes, depth=bs.max_depth, simt=simt, sah=bs.sah_cost)
                r = res[key]
                print(f"{key:40s} {r['msamples']:9.1f} Msamp/s {r['mrays']:9.1f} Mrays/s {r['ms']:8.2f} ms  nodes/ray {r['nodes_per_ray']:.1f} recs/ray {r['recs_per_ray']:.1f}  build {r['build_ms']:.2f} ms depth {r['depth']} sah {r['sah']:.1f}", flush=True)
    print(json.dumps(res))

if __name__ == "__main__":
    main()
