import importlib, sys, torch, numpy as np
sys.path.insert(0, "/root/repo")
rt = importlib.import_module("ray-tracing-cuda_b200")
ctx = rt.Context(0)
scene, cam = rt.scenes.by_name("bunny")
ctx.commit(scene); ctx.set_camera(rt.core.camera_look_at(cam))
H, W = 1080, 1920
q = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
for spp, flags in ((256, 0), (32, rt.abi.RT_RENDER_PROFILE)):
    st = ctx.render_fixed_device(q.data_ptr(), H, W, spp, 0, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR)
    print("spp", spp, "ms", st.render_ms, "seq_fallbacks", st.seq_fallbacks, "max_box_tests_per_ray", st.max_box_tests_per_ray,
          "box/ray", st.node_visits / max(st.rays, 1), "rays", st.rays)
# per-tile cost map at 8 spp with profile: which tiles are pathological?  render tile rows separately is not possible; instead
# time sub-images: crop via camera is not available, so just report NaN radiance pixels
ctx.render_fixed_device(q.data_ptr(), H, W, 8, 0, seed=1, flags=rt.abi.RT_RENDER_CLEAR)
torch.cuda.synchronize()
print("max sum", q.max().item(), "min", q.min().item())
