"""Work-plan sweep (diagnostic, one GPU): kernel time of one rt_render_fixed launch per RT_PLAN value (read per call),
L2 flushed between launches, median of three.  usage: python tools/plan_sweep.py <tag> [scene:H:W:spp ...]"""
import importlib, json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
rt = importlib.import_module("ray-tracing-cuda_b200")

tag = sys.argv[1] if len(sys.argv) > 1 else "x"
cases = sys.argv[2:] or ["cornell_box:1024:1024:128", "cornell_box:1024:1024:1024", "birthday:1080:1920:64", "birthday:1080:1920:512"]
PLANS = {
    128: ["", "16x1:96,4x1", "16x2:96,16x1:16,4x1", "16x4:64,16x1:48,4x1", "16x4:64,16x1:32,8x1:16,4x1", "16x4:96,16x1:16,4x1", "16x3:96,16x1:24,4x1",
          "16x6:96,16x1:24,4x1", "16x4:64,16x2:32,16x1:16,4x1", "16x4:64,16x1:56,4x1", "16x2:64,16x1:48,4x1"],
    1024: ["", "16x9:720,4x1", "16x9:864,16x1:144,4x1", "16x16:768,16x4:128,16x1:96,4x1", "16x16:768,16x1:224,4x1", "16x8:896,16x1:112,4x1",
           "16x16:896,16x2:64,16x1:48,4x1", "16x32:512,16x8:384,16x1:112,4x1", "16x4:960,16x1:48,4x1"],
    64: ["", "4x1", "16x1:48,4x1", "16x2:32,16x1:16,4x1:8,2x1", "16x1:32,8x1:16,4x1"],
    512: ["", "16x4:384,4x1", "16x8:384,16x1:112,4x1", "16x8:448,16x1:48,4x1", "16x4:448,16x1:48,4x1", "16x16:384,16x1:112,4x1"],
}
ctx = rt.Context(0)
flush = torch.empty(256 << 20, dtype=torch.uint8, device="cuda")
out = {}
for case in cases:
    name, H, W, spp = case.split(":")
    H, W, spp = int(H), int(W), int(spp)
    scene, cam = rt.scenes.by_name(name)
    ctx.commit(scene)
    ctx.set_camera(rt.core.camera_look_at(cam))
    sums = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    for plan in PLANS.get(spp, [""]):
        os.environ.pop("RT_PLAN", None)
        if plan:
            os.environ["RT_PLAN"] = plan
        ts = []
        for rep in range(4):
            flush.fill_(1)
            torch.cuda.synchronize()
            st = ctx.render_fixed_device(sums.data_ptr(), H, W, spp, 0, 10, seed=1, flags=rt.abi.RT_RENDER_CLEAR)
            ts.append(st.render_ms)
        ms = sorted(ts[1:])[1]
        out[f"{case} {plan or 'default'}"] = ms
        print(f"{case:28s} {plan or 'default':40s} {ms:9.3f} ms  {H * W * spp / ms / 1e3:9.1f} Msamples/s  items {st.work_items}", flush=True)
os.environ.pop("RT_PLAN", None)
json.dump(out, open(os.path.join(ROOT, "gpurun_out", f"{tag}_plan_sweep.json"), "w"))
