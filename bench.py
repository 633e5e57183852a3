"""bench.py — the reference's headline metric (path Msamples/s) on every BASELINE.json config, on N B200s of one node.

Top level of the JSON line: BASELINE.json configs[1], scenes/cornell_box.cu at 1024x1024, 1024 spp, max depth 10.
`workloads` carries the other configs with the same keys (spheres 400x225x16 at depth 10 and 50, bunny 1080p x256,
birthday 1080p x512, triangle soups 1 M / 10 M at 4K x64), `hit_bit_exact` the primary-ray (t, hit) mismatch counts of
the metric's last clause: against the CPU oracle, against the reference's CUDA code built --fmad=false, and against its
default build (all on full-resolution pixel-centre rays).

A step = one full frame: every rank renders its GetWorkload share of the samples of every pixel (sample partition,
utils.cu:111-113) into exact 64-bit sums and the sums are added at rank 0 with one ncclReduce (replaces D2H +
MPI_Reduce, utils.cu:115-125).  `value` = H*W*spp / step time with the scene resident in HBM; `e2e` = the same frame
through host buffers (scene description H2D + BVH build + render + reduce + resolve + image D2H), i.e. what
Main()/DistributedMain() do per run.

`--impl reference` times the reference's own sources compiled for the host (oracle/_ref, built from /root/reference by
oracle/Makefile) on all host threads, on a bounded sample of the same frame.
`--workload X` benches one config alone (top level = that config); `--only-headline` skips the other configs.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import tempfile
import threading
import time
import zlib

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (scene, H, W, spp, max depth)
    "cornell_box": ("cornell_box", 1024, 1024, 1024, 10),
    "spheres_d10": ("spheres", 225, 400, 16, 10),
    "spheres_d50": ("spheres", 225, 400, 16, 50),
    "bunny": ("bunny", 1080, 1920, 256, 10),
    "birthday": ("birthday", 1080, 1920, 512, 10),
    # BASELINE.json configs[4]: procedural triangle soup, scene data >> L2: HBM-bound roofline
    "soup1m": ("soup1000000", 2160, 3840, 64, 10),
    "soup10m": ("soup10000000", 2160, 3840, 64, 10),
}
WORKLOADS["spheres"] = WORKLOADS["spheres_d10"]
HEADLINE = "cornell_box"
# the other configs in the default run: (timed steps, warm-up steps)
OTHERS = {"spheres_d10": (5, 3), "spheres_d50": (5, 3), "bunny": (3, 3), "birthday": (3, 3), "soup1m": (2, 1), "soup10m": (2, 1)}


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled for the whole run (B200_PROFILING.md); a timed region reads the
    samples that arrived between its start and its end."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "100",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append((time.perf_counter(), ln.strip()))

    def region(self, t0, t1):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        lines = [ln for (t, ln) in list(self.lines) if t0 <= t <= t1 + 0.15]
        if not lines:  # a region shorter than the sampling period: the nearest samples
            lines = [ln for (t, ln) in list(self.lines) if t0 - 0.3 <= t <= t1 + 0.3]
        sm, mx, reasons, power = [], [], set(), []
        for ln in lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                power.append(float(f[3]))
            except ValueError:
                continue
            for name, v in zip(("hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"), f[5:9]):
                if v.lower().startswith("active"):
                    reasons.add(name)
        return {"sm_mhz": float(np.median(sm)) if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "power_w_max": max(power) if power else None, "samples": len(sm), "reasons": sorted(reasons)}

    def stop(self):
        if self.proc:
            self.proc.terminate()
            try:
                self.proc.wait(timeout=5)
            except Exception:
                self.proc.kill()


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def load_profile_constants():
    """Per-launch figures measured with ncu (profiles/): DRAM bytes, lanes per instruction of the dominant kernel."""
    try:
        with open(os.path.join(ROOT, "profiles", "constants.json")) as f:
            return json.load(f)
    except OSError:
        return {}


def scene_of(rt, name):
    return rt.scenes.by_name(name)


def ref_scene_named(ck, name):
    """The reference's own scene (oracle/_ref host build of scenes/<name>.cu) with the same synthetic mesh / texture
    the CUDA arm renders (the model and image files of the reference repo are absent: SURVEY 8d)."""
    pkg = ck.pkg
    if name == "bunny":
        scene, _ = pkg.scenes.by_name(name)
        return ck.RefScene.named(name, 10086, scene.meshes[0], None)
    if name == "birthday":
        scene, _ = pkg.scenes.by_name(name)
        return ck.RefScene.named(name, 10086, None, scene.textures[0])
    return ck.RefScene.named(name, 10086)


def reference_arm(args, workload):
    """The reference's own CPU code for the path (oracle/_ref/libref_host.so) on all host threads."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import checkers as ck

    name, H, W, spp, depth = WORKLOADS[workload]
    if name.startswith("soup"):
        print(json.dumps({"impl": "reference", "unavailable": "no bounded sample of a 1M+ triangle frame finishes on the host "
                          "reference build (2048-triangle leaves, bvh.cuh:71-98)"}))
        return
    base = {"impl": "reference", "metric": "path Msamples/s", "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"scenes/{name}.cu {W}x{H} {spp}spp depth10"}}
    kind = "reference" if ck.have_ref() else "port"
    threads = os.cpu_count() or 1
    if kind == "reference":
        ref = ref_scene_named(ck, name)
        run = lambda rows, s: ref.render(rows, W, s, post=0, nthreads=threads)  # noqa: E731
    else:
        pkg = ck.pkg
        scene, cam = pkg.scenes.by_name(name)
        o = ck.OracleScene(scene)
        o.camera_look_at(cam)
        run = lambda rows, s: o.render_reference(rows, W, s, 10086, nthreads=threads)  # noqa: E731
    # bounded sample: the full frame at spp_s samples per pixel, sized from a probe to ~6 s per step
    t0 = time.perf_counter()
    run(H // 16, 1)
    probe = time.perf_counter() - t0
    full_1spp = probe * 16
    spp_s = max(1, min(spp, int(6.0 / max(full_1spp, 1e-6))))
    for _ in range(min(args.warmup, 1)):
        run(H, spp_s)
    times = []
    for _ in range(args.steps):
        t0 = time.perf_counter()
        run(H, spp_s)
        times.append(time.perf_counter() - t0)
    t = float(np.mean(times))
    v = H * W * spp_s / t / 1e6
    base.update({"value": v, "ms_per_step": t * 1e3,
                 "cpu_baseline": {"value": v, "unit": "Msamples/s", "cores": threads, "kind": kind,
                                  "sample": f"full {W}x{H} frame at {spp_s} spp per step (of {spp})"},
                 "e2e": {"value": v, "unit": "Msamples/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
                 "gpu_launches": 0})
    print(json.dumps(base))


def cpu_baseline(name, H, W, spp, budget_s=10.0):
    import checkers as ck

    threads = os.cpu_count() or 1
    if ck.have_ref():
        kind = "reference"
        ref = ref_scene_named(ck, name)
        run = lambda s, h, w: ref.render(h, w, s, post=0, nthreads=threads)  # noqa: E731
    else:
        kind = "port"
        scene, cam = ck.pkg.scenes.by_name(name)
        o = ck.OracleScene(scene)
        o.camera_look_at(cam)
        run = lambda s, h, w: o.render_reference(h, w, s, 10086, nthreads=threads)  # noqa: E731
    # bounded sample: probe a quarter-resolution frame (same camera, same field of view); if a full frame at 1 spp
    # would not fit the budget (bunny: 2048-triangle leaves), the quarter-resolution frame is the sample
    h4, w4 = max(H // 4, 1), max(W // 4, 1)
    t0 = time.perf_counter()
    run(1, h4, w4)
    tp = time.perf_counter() - t0
    if tp * 16 > budget_s:
        hs, ws, label = h4, w4, f"quarter-resolution {w4}x{h4} frame"
        spp_s = max(1, min(spp, int(budget_s / max(tp, 1e-6))))
    else:
        hs, ws, label = H, W, f"full {W}x{H} frame"
        spp_s = max(1, min(spp, int(budget_s / max(tp * 16, 1e-6))))
    t0 = time.perf_counter()
    run(spp_s, hs, ws)
    t = time.perf_counter() - t0
    return {"value": hs * ws * spp_s / t / 1e6, "unit": "Msamples/s", "cores": threads, "kind": kind,
            "sample": f"{label} at {spp_s} spp (of {spp}), {t:.1f} s"}


def _ref_cuda_args(name, scene, tmp):
    """Command-line extras that hand the reference CUDA harness the same synthetic mesh / texture."""
    extra, ref_scene = [], name
    if name == "bunny":
        fp = os.path.join(tmp, "faces.bin")
        scene.meshes[0].astype(np.float32).tofile(fp)
        extra = ["--faces", fp]
    if name == "birthday":
        fp = os.path.join(tmp, "tex.bin")
        np.ascontiguousarray(scene.textures[0]).tofile(fp)
        extra = ["--tex", fp, str(scene.textures[0].shape[0]), str(scene.textures[0].shape[1])]
    return ref_scene, extra


def ref_cuda_baseline(name, H, W, scene, spp=16):
    """The reference's own CUDA build on this GPU (baseline/_ref/ref_cuda, built from /root/reference by
    baseline/Makefile): its PathTracing kernel timed its own way at `spp` samples (the kernel is linear in spp)."""
    exe = os.path.join(ROOT, "baseline", "_ref", "ref_cuda")
    if not os.path.exists(exe):
        return {"unavailable": "baseline/_ref/ref_cuda not built"}
    tmp = tempfile.mkdtemp()
    ref_scene, extra = _ref_cuda_args(name, scene, tmp)
    if name == "bunny":
        spp = 2
    try:
        p = subprocess.run([exe, ref_scene, str(H), str(W), str(spp), "3"] + extra, capture_output=True, text=True, timeout=240)
        r = json.loads(p.stdout.strip().splitlines()[-1])
        return {"value": r["msamples_per_s"], "unit": "Msamples/s", "kind": "reference CUDA build, sm_100a, --fmad default",
                "sample": f"full {W}x{H} frame at {spp} spp, best of 3, kernel-only (cudaEvents) as utils.cu:155-170", "ms": r["ms_best"]}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)[:200]}


REF_HIT = np.dtype([("t", "<f8"), ("u", "<f8"), ("v", "<f8"), ("normal", "<f4", 3), ("hit", "<i4")])


def hit_bit_exact(rt, ctx, name, H, W, scene, camspec, oracle_budget_s=8.0):
    """BASELINE.json's `hit bit-exactness`: closest hits of full-resolution pixel-centre primary rays through
    rt_trace_closest against (a) the CPU oracle (t bits, object / face ids, normal), (b) world->Hit of the reference's
    own CUDA code built --fmad=false and (c) its default build (t bits, hit flag, normal; the reference has no ids)."""
    import checkers as ck

    out = {"rays": H * W}
    scene_b, cam_b = scene, camspec
    if name == "spheres":
        # the layout is drawn inside user code (spheres.cu:54-80): nvcc's device front end evaluates the draws of one
        # argument list left to right, g++ (the host build the oracle is pinned to) right to left; the GPU reference
        # therefore renders the "ltr" scene, and so do we here
        scene_b, cam_b, _ = rt.scenes.spheres(arg_order="ltr")
    camera = rt.core.camera_look_at(cam_b)
    # (a) oracle: the reference's semantics on the host (its own BVH, AABB::Hit included); a stride keeps it in budget.
    # The oracle is pinned to the reference's HOST build, whose pow() is glibc's (correctly rounded); the reference's
    # CUDA build and the product call CUDA's pow, 1 ulp off on 6-14 % of operands (pow_kat_device).  The comparison
    # with the oracle therefore commits the scene with RT_SPHERE_POW=ieee; `mismatch_vs_oracle_cuda_pow` is the count
    # in the product's default mode (sphere hits only).
    os.environ["RT_SPHERE_POW"] = "ieee"
    try:
        ctx.commit(scene_b)
        ctx.set_camera(camera)
        rays = ctx.primary_rays(H, W)
        mine_ieee = ctx.trace_closest(rays)
    finally:
        del os.environ["RT_SPHERE_POW"]
    ctx.commit(scene_b)
    ctx.set_camera(camera)
    mine = ctx.trace_closest(rays)
    o = ck.OracleScene(scene_b, camera)
    stride = 1
    t0 = time.perf_counter()
    o.trace(rays[::64], quirks=1)
    dt = (time.perf_counter() - t0) * 64
    while dt / stride > oracle_budget_s:
        stride *= 2
    want = o.trace(rays[::stride], quirks=1) if stride > 1 else o.trace(rays, quirks=1)
    fields = ("t", "obj", "sub", "normal", "material")
    out["mismatch_vs_oracle"] = int(len(ck.hits_equal(mine_ieee[::stride], want, fields)))
    out["mismatch_vs_oracle_cuda_pow"] = int(len(ck.hits_equal(mine[::stride], want, fields)))
    out["oracle_rays"] = int(len(want))
    out["oracle_note"] = "oracle = reference semantics on the host; product committed with RT_SPHERE_POW=ieee for this column"
    # (b), (c) the reference's CUDA code
    tmp = tempfile.mkdtemp()
    ref_scene, extra = _ref_cuda_args(name, scene_b, tmp)
    rp = os.path.join(tmp, "rays.bin")
    rays.tofile(rp)
    hit_m = mine["obj"] >= 0
    for key, exe in (("ref_cuda_nofmad", "ref_cuda_nofmad"), ("ref_cuda_default", "ref_cuda")):
        path = os.path.join(ROOT, "baseline", "_ref", exe)
        if not os.path.exists(path):
            out[f"mismatch_vs_{key}"] = None
            continue
        hp = os.path.join(tmp, key + "_hits.bin")
        try:
            subprocess.run([path, ref_scene, str(H), str(W), "1", "0"] + extra + ["--rays", rp, hp], capture_output=True,
                           text=True, timeout=240, check=True)
            rh = np.fromfile(hp, dtype=REF_HIT)
            ref_hit = rh["hit"] != 0
            both = ref_hit & hit_m
            tb = (rh["t"].view(np.uint64) != mine["t"].view(np.uint64)) & both
            nb = (rh["normal"].view(np.uint32) != mine["normal"].view(np.uint32)).any(axis=1) & both & (mine["t"] < 1e9)
            flag = ref_hit != hit_m
            rel = np.abs(rh["t"][both] - mine["t"][both]) / np.maximum(np.abs(mine["t"][both]), 1e-30)
            out[f"mismatch_vs_{key}"] = int((tb | nb | flag).sum())
            out[f"{key}_detail"] = {"t_bits": int(tb.sum()), "normal_bits": int(nb.sum()), "hit_flag": int(flag.sum()),
                                    "max_rel_t": float(rel.max() if len(rel) else 0.0)}
        except Exception as e:  # noqa: BLE001
            out[f"mismatch_vs_{key}"] = None
            out[f"{key}_detail"] = {"error": repr(e)[:160]}
    d = out.get("ref_cuda_nofmad_detail", {})
    out["max_rel_t"] = d.get("max_rel_t")
    return out


def pow_kat():
    exe = os.path.join(ROOT, "baseline", "_ref", "ref_cuda_nofmad")
    if not os.path.exists(exe):
        return None
    try:
        p = subprocess.run([exe, "powkat", "0", "0", "0", "0"], capture_output=True, text=True, timeout=120)
        return json.loads(p.stdout.strip().splitlines()[-1])
    except Exception as e:  # noqa: BLE001
        return {"error": repr(e)[:160]}


class DeviceArray:
    """A raw device address as something torch.as_tensor can wrap without a copy."""

    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


class Runner:
    """One context per process; benches one workload at a time."""

    def __init__(self, args):
        import torch
        import torch.distributed as dist

        self.torch, self.dist, self.args = torch, dist, args
        self.rt = importlib.import_module("ray-tracing-cuda_b200")
        self.rdist = importlib.import_module("ray-tracing-cuda_b200.distributed")
        self.rank, self.world, self.local = self.rdist.init_process_group()
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        self.ctx = self.rt.Context(self.local)
        if self.world > 1:
            self.rdist.bind_nccl(self.ctx)
        self.stream = torch.cuda.current_stream()
        self.flush = torch.empty(256 << 20, dtype=torch.uint8, device=self.dev)  # > 126 MB L2
        self.peaks, self.peak_kind = load_peaks()
        self.consts = load_profile_constants()
        self.props = torch.cuda.get_device_properties(self.local)
        self.sampler = ClockSampler(self.local)
        if self.rank == 0:
            self.sampler.start()

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def timed(self, fn, steps, warmup):
        torch = self.torch
        for _ in range(warmup):
            fn()
            self.flush.fill_(1)
        times = []
        self.barrier()
        for _ in range(steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            self.barrier()
            e0.record(self.stream)
            fn()
            e1.record(self.stream)
            self.barrier()
            times.append(e0.elapsed_time(e1))
            self.flush.fill_(1)  # L2 flush between timed iterations
        self.barrier()
        t = torch.tensor([float(np.sum(times))], device=self.dev)
        if self.world > 1:
            self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX)
        return float(t.item()) / steps

    def bench(self, workload, steps, warmup, baselines):
        torch, rt, ctx, args = self.torch, self.rt, self.ctx, self.args
        rank, world = self.rank, self.world
        name, H, W, spp, depth = WORKLOADS[workload]
        if args.spp:
            spp = args.spp
        if args.depth > 0:
            depth = args.depth
        scene, camspec = scene_of(rt, name)
        camera = rt.core.camera_look_at(camspec)
        build = ctx.commit(scene)
        ctx.set_camera(camera)
        count, begin = self.rdist.partition(spp, rank, world)
        flags = {"auto": 0, "fused": rt.abi.RT_RENDER_FUSED, "wavefront": rt.abi.RT_RENDER_WAVEFRONT,
                 "wavefront_raysort": rt.abi.RT_RENDER_WAVEFRONT | rt.abi.RT_RENDER_RAYSORT}[args.strategy]
        sp = self.stream.cuda_stream
        sums = torch.zeros(H, W, 3, device=self.dev, dtype=torch.int64)
        rgb = torch.zeros(H, W, 3, device=self.dev)
        rgb_host = torch.zeros(H, W, 3).pin_memory()
        CLEAR = rt.abi.RT_RENDER_CLEAR
        # N > 1: the sums of all ranks are added, resolved and gathered at rank 0 by one kernel per rank over NVLink
        # peer memory (rt_reduce_resolve_peers; --reduce nccl = one ncclReduce of the sums + a resolve at rank 0).
        peers = world > 1 and args.reduce == "peers"
        sums_ptr = sums.data_ptr()
        if peers:
            sums_ptr, _, rgb_ptr = self.rdist.open_peer_frame(ctx, H, W)
            rgb = torch.as_tensor(DeviceArray(rgb_ptr, (H, W, 3), "<f4"), device=self.dev)

        def reduce_to_root(resolve):
            if peers:
                ctx.reduce_resolve_peers(spp, 0, sp)
            else:
                if world > 1:
                    ctx.reduce_fixed(None, sums_ptr, H * W * 3, 0, sp)
                if resolve and rank == 0:
                    ctx.resolve_fixed_device(sums_ptr, H, W, spp, None, rgb.data_ptr(), None, sp)

        def step_device():
            ctx.render_fixed_device(sums_ptr, H, W, count, begin, depth, seed=1, flags=flags | CLEAR, stream=sp)
            if world > 1:
                reduce_to_root(False)

        def step_e2e():
            ctx.commit(scene)  # host description -> device records + BVH (H2D inside)
            ctx.set_camera(camera)
            ctx.render_fixed_device(sums_ptr, H, W, count, begin, depth, seed=1, flags=flags | CLEAR, stream=sp)
            reduce_to_root(True)
            if rank == 0:
                rgb_host.copy_(rgb, non_blocking=True)
            self.stream.synchronize()

        t0 = time.perf_counter()
        ms = self.timed(step_device, steps, warmup)
        t1 = time.perf_counter()
        clocks = self.sampler.region(t0, t1) if rank == 0 else None
        # one more step for ray counts and the kernel-only time, then a stats launch (counters of the same frame at
        # reduced spp: per-ray figures are spp independent)
        st = ctx.render_fixed_device(sums_ptr, H, W, count, begin, depth, seed=1, flags=flags | CLEAR, stream=sp)
        prof_spp = max(1, min(count, 4))
        st_prof = ctx.render_fixed_device(sums_ptr, H, W, prof_spp, begin, depth, seed=1,
                                          flags=flags | CLEAR | rt.abi.RT_RENDER_PROFILE, stream=sp)
        ms_e2e = self.timed(step_e2e, max(1, min(steps, 3)), 1)
        if peers:
            ctx.peer_frame_check(sp)  # raises if a rank timed out inside the reduce
            image_crc = int(zlib.crc32(rgb_host.numpy().tobytes())) if rank == 0 else 0
            del rgb
            self.barrier()
            ctx.peer_frame_destroy()
        else:
            image_crc = int(zlib.crc32(rgb_host.numpy().tobytes())) if rank == 0 else 0
        rays = torch.tensor([float(st.rays), float(st.rays_ref_equiv)], device=self.dev, dtype=torch.float64)
        if world > 1:
            self.dist.all_reduce(rays)
        if rank != 0:
            return None
        peaks, consts = self.peaks, self.consts
        samples = H * W * spp
        h2d = int(scene.hitables.nbytes + scene.materials.nbytes + sum(m.nbytes for m in scene.meshes) +
                  sum(t.nbytes for t in scene.textures) + 96)
        kern_ms = st.render_ms
        fused_run = st.iterations == 0
        sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
        # Roofline of the dominant kernel.  Cache-resident scenes (C1-C4): the bound is SM instruction issue (SURVEY 8d,
        # DESIGN 3): peak = SMs x 4 schedulers x f_clk warp-instructions/s at the clock sampled DURING the run.
        # Algorithmic work per launch (SURVEY 8d figures x the counters of a stats launch of the same frame): 23 flops
        # per box test, 52 per ray-triangle record test, 23 per accepted hit, /1.3 (FMA-fused instruction count) + 12
        # control instructions per box test; one warp instruction = 32 lanes.
        boxes, recs_tested = float(st_prof.node_visits), float(st_prof.prim_tests)
        scale = float(count) / max(prof_spp, 1)
        thread_inst = ((23.0 * boxes + 52.0 * recs_tested + 23.0 * float(st_prof.rays)) / 1.3 + 12.0 * boxes) * scale
        issue_peak = self.props.multi_processor_count * 4 * sm_mhz * 1e6
        achieved = thread_inst / 32.0 / (kern_ms * 1e-3)
        is_soup = name.startswith("soup")
        base_name = workload
        traffic = consts.get("dram_bytes_per_launch", {}).get(base_name)
        if traffic is not None and is_soup:  # soups: node / record misses scale with the samples of the launch
            traffic = int(traffic * float(count) / consts.get("dram_bytes_capture_spp", {}).get(base_name, count))
        roofline = {"bound": "issue", "achieved": achieved / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                    "frac": achieved / issue_peak, "traffic": traffic,
                    "traffic_source": "dram__bytes_read+write of the committed ncu capture of this workload (profiles/constants.json"
                                      f", captured at {consts.get('dram_bytes_capture_spp', {}).get(base_name)} spp"
                                      + (", scaled to this launch's samples" if is_soup else "") + "); not measured in this run",
                    "kernel": "k_render" if fused_run else "k_wf_extend", "kernel_ms": kern_ms, "sm_mhz": sm_mhz,
                    "peak_source": "148 SMs x 4 issue slots x sampled SM clock (B200_PROFILING.md / SURVEY 8d); not a tensor or HBM bound",
                    "algorithmic": {"box_tests_per_ray": boxes / max(float(st_prof.rays), 1.0),
                                    "record_tests_per_ray": recs_tested / max(float(st_prof.rays), 1.0),
                                    "thread_inst_per_ray": thread_inst / max(float(st.rays), 1.0)},
                    "measured_ncu": consts.get("ncu", {}).get(base_name)}
        if is_soup:
            # Scene data (48 B records + 32 B nodes) is far larger than L2 and bounce rays are incoherent: the bound is
            # HBM.  Algorithmic bytes per launch = every box test reads its 32-byte node, every record test its 48-byte
            # record, plus the sums read + written once.  `frac` rises with the box tests of a worse tree,
            # so the line also carries `dram_frac`: the DRAM bytes ncu measured for this workload / kernel time / peak.
            alg = (32.0 * boxes + 48.0 * recs_tested) * scale + H * W * 48
            a_gbs = alg / (kern_ms * 1e-3) / 1e9
            roofline.update({"bound": "hbm", "achieved": a_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                             "frac": a_gbs / peaks["hbm_gbs"], "peak_source": f"{self.peak_kind} MEASURED_PEAKS.json hbm_gbs",
                             "dram_frac": (traffic / (kern_ms * 1e-3) / 1e9 / peaks["hbm_gbs"]) if traffic else None,
                             "issue": {"achieved": achieved / 1e9, "peak": issue_peak / 1e9, "frac": achieved / issue_peak}})
        label = (f"synthetic triangle soup {build.n_tris} tris {W}x{H} {spp}spp depth{depth}" if is_soup
                 else f"scenes/{name}.cu {W}x{H} {spp}spp depth{depth}")
        out = {"metric": "path Msamples/s", "value": samples / ms / 1e3, "unit": "Msamples/s", "n_gpus": world, "steps": steps,
               "warmup": warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
               "dtype": "f32", "data": "synthetic",
               "config": {"workload": label, "partition": f"spp/{world}",
                          "reduce": ("nvlink peer frames, one kernel per rank (sum + resolve + gather)" if peers else
                                     "ncclReduce int64 + resolve at rank 0") if world > 1 else None,
                          "image_crc32": image_crc,
                          "l2": "flushed between timed steps (256 MiB write)", "strategy": args.strategy,
                          "prims": build.n_prims, "nodes": build.n_nodes,
                          "chunk_samples": int(st.chunk_samples), "work_items": int(st.work_items)},
               "mrays_per_s": float(rays[0].item()) / ms / 1e3, "mrays_ref_equiv_per_s": float(rays[1].item()) / ms / 1e3,
               "rays_per_sample": float(rays[0].item()) / samples, "frame_ms": ms,
               "e2e": {"value": samples / ms_e2e / 1e3, "unit": "Msamples/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d,
                       "d2h_bytes_per_step": H * W * 12},
               "gpu_launches": int((st.kernel_launches + (3 if peers else 0)) * steps),
               "roofline": roofline, "clocks": clocks,
               "build": {"ms": build.build_ms, "max_depth": build.max_depth, "sah": build.sah_cost,
                         "tree": "ploc" if build.reserved0_ > 0 else "lbvh", "ploc_rounds": abs(build.reserved0_),
                         "sah_other_tree": float(np.array([build.reserved1_], np.int32).view(np.float32)[0]) if build.reserved0_ else None}}
        if baselines and is_soup:
            out["cpu_baseline"] = {"unavailable": "the host reference build scans meshes with 2048-triangle leaves (bvh.cuh:71-98): "
                                                  "no bounded sample of a 1M+ triangle frame finishes in the bench budget"}
        elif baselines:
            if workload != "spheres_d50":  # the reference hard-codes depth 10 (ray_tracing.cu:10)
                out["cpu_baseline"] = cpu_baseline(name, H, W, spp, 10.0 if workload == HEADLINE else 4.0)
                out["ref_cuda_baseline"] = ref_cuda_baseline(name, H, W, scene)
            if workload != "spheres_d10" and workload != "spheres_d50" or workload == "spheres_d10":
                pass
        return out


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default=None, choices=sorted(WORKLOADS), help="bench this config alone")
    ap.add_argument("--only-headline", action="store_true", help="skip the other BASELINE configs and hit_bit_exact")
    ap.add_argument("--spp", type=int, default=0, help="override spp (diagnostic; the headline uses the config's)")
    ap.add_argument("--depth", type=int, default=0, help="override the bounce limit (diagnostic)")
    ap.add_argument("--strategy", default="auto", choices=["auto", "fused", "wavefront", "wavefront_raysort"])
    ap.add_argument("--reduce", default="peers", choices=["peers", "nccl"],
                    help="N > 1: one kernel over NVLink peer frames (default) or ncclReduce + resolve at rank 0")
    ap.add_argument("--no-cpu-baseline", action="store_true")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return reference_arm(args, args.workload or HEADLINE)

    r = Runner(args)
    single = args.workload is not None
    head = args.workload or HEADLINE
    baselines = (not args.no_cpu_baseline) and r.world == 1
    out = r.bench(head, args.steps, args.warmup, baselines)
    if not single and not args.only_headline and not args.spp and not args.depth:
        others = {}
        for w, (steps, warm) in OTHERS.items():
            res = r.bench(w, steps, warm, baselines)
            if res is not None:
                for k in ("metric", "unit", "higher_is_better", "scaling", "vs_baseline", "dtype", "data", "n_gpus"):
                    res.pop(k, None)
                others[w] = res
        if r.rank == 0:
            out["workloads"] = others
            if baselines:
                hb = {}
                for w in ("cornell_box", "spheres_d10", "bunny", "birthday"):
                    name, H, W, _, _ = WORKLOADS[w]
                    scene, camspec = scene_of(r.rt, name)
                    hb[name] = hit_bit_exact(r.rt, r.ctx, name, H, W, scene, camspec)
                hb["pow_kat_device"] = pow_kat()
                out["hit_bit_exact"] = hb
    if r.rank == 0:
        r.sampler.stop()
        print(json.dumps(out))
    if r.world > 1:
        r.dist.destroy_process_group()


if __name__ == "__main__":
    main()
