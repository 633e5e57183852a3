"""bench.py — the reference's headline metric (path Msamples/s) on BASELINE.json configs[1]:
scenes/cornell_box.cu at 1024x1024, 1024 spp, max depth 10, on N B200s of one node.

A step = one full frame: every rank renders its GetWorkload share of the 1024 samples per pixel
(sample partition, utils.cu:111-113) and the float accumulation buffers are summed at rank 0 with
one ncclReduce (replaces D2H + MPI_Reduce, utils.cu:115-125).  `value` = H*W*spp / step time with
the scene resident in HBM; `e2e` = the same frame through host buffers (scene description H2D +
BVH build + render + reduce + resolve + image D2H), i.e. what Main()/DistributedMain() do per run.

`--impl reference` times the reference's own sources compiled for the host (oracle/_ref, built from
/root/reference by oracle/Makefile) on all host threads, on a bounded sample of the same frame.
"""
import argparse
import importlib
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

WORKLOADS = {
    # name: (scene, H, W, spp)
    "cornell_box": ("cornell_box", 1024, 1024, 1024),
    "spheres": ("spheres", 225, 400, 16),
    "bunny": ("bunny", 1080, 1920, 256),
    "birthday": ("birthday", 1080, 1920, 512),
    # BASELINE.json configs[4]: procedural triangle soup, scene data >> L2: HBM-bound roofline
    "soup1m": ("soup1000000", 2160, 3840, 64),
    "soup10m": ("soup10000000", 2160, 3840, 64),
}
MAX_DEPTH = 10


class ClockSampler:
    """nvidia-smi clocks / throttle reasons sampled DURING the timed region (B200_PROFILING.md)."""

    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,clocks_event_reasons.hw_slowdown,"
         "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu):
        self.gpu, self.proc, self.lines = gpu, None, []

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits", "-lms", "200",
                                          "-i", str(self.gpu)], stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.t = threading.Thread(target=self._read, daemon=True)
            self.t.start()
        except OSError:
            self.proc = None

    def _read(self):
        for ln in self.proc.stdout:
            self.lines.append(ln.strip())

    def stop(self):
        if not self.proc:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, power = [], [], set(), []
        for ln in self.lines:
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


def load_peaks():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return json.load(f), "measured"
    except OSError:
        return {"hbm_gbs": 6650.0, "sm_max_mhz": 1965.0}, "fallback"


def load_profile_constants():
    """Per-sample figures measured with ncu (profiles/): warp instructions per sample of the dominant kernel."""
    try:
        with open(os.path.join(ROOT, "profiles", "constants.json")) as f:
            return json.load(f)
    except OSError:
        return {}


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

    name, H, W, spp = WORKLOADS[workload]
    if name.startswith("soup"):
        print(json.dumps({"impl": "reference", "unavailable": "no bounded sample of a 1M+ triangle frame finishes on the host "
                          "reference build (2048-triangle leaves, bvh.cuh:71-98)"}))
        return
    base = {"impl": "reference", "metric": "path Msamples/s", "unit": "Msamples/s", "n_gpus": args.gpus,
            "steps": args.steps, "warmup": args.warmup, "higher_is_better": True, "scaling": "strong",
            "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": {"workload": f"scenes/{name}.cu {W}x{H} {spp}spp depth{MAX_DEPTH}"}}
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
    # bounded sample: the top `rows` rows of the frame at 1 spp, sized from a probe to ~6 s per step
    # (the host harness renders rows [0, rows) of a rows x W image: same camera rays per row only if
    #  rows == H, so the sample keeps rows == H and bounds spp instead)
    t0 = time.perf_counter()
    run(H // 16, 1) if kind == "port" else ref.render(H // 16, W, 1, post=0, nthreads=threads)
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


def cpu_baseline(name, H, W, spp, budget_s=12.0):
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


def ref_cuda_baseline(name, H, W, scene, spp=16):
    """The reference's own CUDA build on this GPU (baseline/_ref/ref_cuda, built from /root/reference by
    baseline/Makefile): its PathTracing kernel timed its own way at `spp` samples (the kernel is linear in spp)."""
    import tempfile
    exe = os.path.join(ROOT, "baseline", "_ref", "ref_cuda")
    if not os.path.exists(exe):
        return {"unavailable": "baseline/_ref/ref_cuda not built"}
    extra = []
    tmp = tempfile.mkdtemp()
    if name == "bunny":
        fp = os.path.join(tmp, "faces.bin")
        scene.meshes[0].astype(np.float32).tofile(fp)
        extra = ["--faces", fp]
        spp = 2
    if name == "birthday":
        fp = os.path.join(tmp, "tex.bin")
        np.ascontiguousarray(scene.textures[0]).tofile(fp)
        extra = ["--tex", fp, str(scene.textures[0].shape[0]), str(scene.textures[0].shape[1])]
    try:
        p = subprocess.run([exe, name, str(H), str(W), str(spp), "3"] + extra, capture_output=True, text=True, timeout=240)
        r = json.loads(p.stdout.strip().splitlines()[-1])
        return {"value": r["msamples_per_s"], "unit": "Msamples/s", "kind": "reference CUDA build, sm_100a, --fmad default",
                "sample": f"full {W}x{H} frame at {spp} spp, best of 3, kernel-only (cudaEvents) as utils.cu:155-170", "ms": r["ms_best"]}
    except Exception as e:  # noqa: BLE001
        return {"unavailable": repr(e)[:200]}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--workload", default="cornell_box", choices=sorted(WORKLOADS))
    ap.add_argument("--spp", type=int, default=0, help="override spp (diagnostic; the headline uses the config's)")
    ap.add_argument("--depth", type=int, default=0,
                    help="bounce limit (default 10 = the reference's TRACE_DEPTH_LIMIT, ray_tracing.cu:10; BASELINE config 1 "
                         "also names depth 50: --workload spheres --depth 50)")
    ap.add_argument("--strategy", default="auto", choices=["auto", "fused", "wavefront"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-split", action="store_true", help="do not pass RT_RENDER_SPLIT (diagnostic)")
    ap.add_argument("--ordered", action="store_true",
                    help="pass RT_RENDER_ORDERED (experimental: heaviest tile first on meshes; off by default)")
    args = ap.parse_args()
    args.warmup = max(args.warmup, 0)
    if args.impl == "reference":
        return reference_arm(args, args.workload)

    import torch
    import torch.distributed as dist

    rt = importlib.import_module("ray-tracing-cuda_b200")
    from importlib import import_module
    rdist = import_module("ray-tracing-cuda_b200.distributed")

    rank, world, local = rdist.init_process_group()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    name, H, W, spp = WORKLOADS[args.workload]
    if args.spp:
        spp = args.spp
    MAX_DEPTH = args.depth if args.depth > 0 else globals()["MAX_DEPTH"]
    scene, camspec = rt.scenes.by_name(name)
    ctx = rt.Context(local)
    camera = rt.core.camera_look_at(camspec)
    build = ctx.commit(scene)
    ctx.set_camera(camera)
    if world > 1:
        rdist.bind_nccl(ctx)
    count, begin = rdist.partition(spp, rank, world)
    flags = {"auto": 0, "fused": rt.abi.RT_RENDER_FUSED, "wavefront": rt.abi.RT_RENDER_WAVEFRONT}[args.strategy]
    stream = torch.cuda.current_stream()
    sp = stream.cuda_stream
    accum = torch.zeros(H, W, 3, device=dev)
    rgb = torch.zeros(H, W, 3, device=dev)
    rgb_host = torch.zeros(H, W, 3).pin_memory()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)  # > 126 MB L2

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def step_device():
        ctx.render_device(accum.data_ptr(), H, W, count, begin, MAX_DEPTH, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR,
                          stream=sp)
        if world > 1:
            ctx.reduce(None, accum.data_ptr(), H * W * 3, 0, sp)

    def step_e2e():
        ctx.commit(scene)  # host description -> device records + BVH (H2D inside)
        ctx.set_camera(camera)
        ctx.render_device(accum.data_ptr(), H, W, count, begin, MAX_DEPTH, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR,
                          stream=sp)
        if world > 1:
            ctx.reduce(None, accum.data_ptr(), H * W * 3, 0, sp)
        if rank == 0:
            ctx.resolve_device(accum.data_ptr(), H, W, spp, rgb.data_ptr(), None, sp)
            rgb_host.copy_(rgb, non_blocking=True)
        stream.synchronize()

    def timed(fn, steps, warmup):
        for _ in range(warmup):
            fn()
            flush.fill_(1)
        times = []
        barrier()
        for _ in range(steps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            barrier()
            e0.record(stream)
            fn()
            e1.record(stream)
            barrier()
            times.append(e0.elapsed_time(e1))
            flush.fill_(1)  # L2 flush between timed iterations
        barrier()
        t = torch.tensor([float(np.sum(times))], device=dev)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item()) / steps

    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(step_device, args.steps, args.warmup)
    clocks = sampler.stop() if rank == 0 else None
    # one more step with stats for ray counts and the kernel-only time
    st = ctx.render_device(accum.data_ptr(), H, W, count, begin, MAX_DEPTH, seed=1, flags=flags | rt.abi.RT_RENDER_CLEAR,
                           stream=sp)
    # stats launch (counters of the same frame at reduced spp: per-ray figures are spp independent)
    prof_spp = max(1, min(count, 4))
    st_prof = ctx.render_device(accum.data_ptr(), H, W, prof_spp, begin, MAX_DEPTH, seed=1,
                                flags=flags | rt.abi.RT_RENDER_CLEAR | rt.abi.RT_RENDER_PROFILE, stream=sp)
    ms_e2e = timed(step_e2e, max(1, min(args.steps, 3)), 1)

    rays = torch.tensor([float(st.rays), float(st.rays_ref_equiv)], device=dev, dtype=torch.float64)
    if world > 1:
        dist.all_reduce(rays)
    if rank != 0:
        if world > 1:
            dist.destroy_process_group()
        return
    peaks, peak_kind = load_peaks()
    consts = load_profile_constants()
    samples = H * W * spp
    value = samples / ms / 1e3
    h2d = int(scene.hitables.nbytes + scene.materials.nbytes + sum(m.nbytes for m in scene.meshes) +
              sum(t.nbytes for t in scene.textures) + 96)
    kern_ms = st.render_ms
    fused_run = st.iterations == 0
    kernel = "k_render" if fused_run else "k_wf_extend"
    sm_mhz = (clocks or {}).get("sm_mhz") or peaks.get("sm_max_mhz", 1965.0)
    props = torch.cuda.get_device_properties(local)
    # Roofline of the dominant kernel.  The scene (nodes + records) is L1/L2 resident, so the bound is SM
    # instruction issue: peak = SMs x 4 schedulers x f_clk warp-instructions/s at the
    # clock sampled DURING the run.  Algorithmic work per launch (SURVEY §8d figures x the counters of a stats
    # launch of the same frame): 23 flops per box test, 52 per ray-triangle record test, 23 per accepted hit,
    # /1.3 (FMA-fused instruction count) + 12 control instructions per box test; one warp instruction = 32 lanes.
    boxes, recs_tested = float(st_prof.node_visits), float(st_prof.prim_tests)
    scale = float(count) / max(prof_spp, 1)  # the stats launch renders prof_spp samples per pixel
    thread_inst = ((23.0 * boxes + 52.0 * recs_tested + 23.0 * float(st_prof.rays)) / 1.3 + 12.0 * boxes) * scale
    issue_peak = props.multi_processor_count * 4 * sm_mhz * 1e6
    achieved = thread_inst / 32.0 / (kern_ms * 1e-3)
    alg_bytes = H * W * 24 + build.n_prims * 48 + build.n_nodes * 32  # accumulation r+w, scene read once
    hbm_achieved = alg_bytes / (kern_ms * 1e-3) / 1e9
    is_soup = name.startswith("soup")
    # DRAM bytes per launch from the committed ncu capture of the same workload (profiles/constants.json).  Cache-
    # resident scenes: the traffic is the accumulation buffer + one read of the scene, independent of spp.  Soups:
    # node/record misses dominate and scale with the samples of the launch (capture spp -> this launch's spp).
    traffic = consts.get("dram_bytes_per_launch", {}).get(args.workload)
    if traffic is not None and is_soup:
        traffic = int(traffic * float(count) / consts.get("dram_bytes_capture_spp", {}).get(args.workload, count))
    roofline = {"bound": "issue", "achieved": achieved / 1e9, "peak": issue_peak / 1e9, "unit": "Gwarp-inst/s",
                "frac": achieved / issue_peak, "traffic": traffic,
                "kernel": kernel, "kernel_ms": kern_ms, "sm_mhz": sm_mhz,
                "peak_source": "148 SMs x 4 issue slots x sampled SM clock (B200_PROFILING.md / SURVEY 8d); not a tensor or HBM bound",
                "algorithmic": {"box_tests_per_ray": boxes / max(float(st_prof.rays), 1.0),
                                "record_tests_per_ray": recs_tested / max(float(st_prof.rays), 1.0),
                                "thread_inst_per_ray": thread_inst / max(float(st.rays), 1.0)},
                "hbm": {"bound": "hbm", "achieved": hbm_achieved, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                        "frac": hbm_achieved / peaks["hbm_gbs"], "peak_source": f"{peak_kind} MEASURED_PEAKS.json hbm_gbs",
                        "note": "scene is cache resident: HBM is not the binding resource"},
                "measured_ncu": consts.get("ncu", {}).get(args.workload)}
    if is_soup:
        # Scene data (48 B records + 32 B nodes) is far larger than L2 and bounce rays are incoherent: the bound is
        # HBM.  Algorithmic bytes per launch = every node visit reads its 32-byte node, every record test its 48-byte
        # record, plus the accumulation buffer read + written once; cache hits make the measured DRAM
        # traffic (profiles/, `traffic`) smaller than this.
        alg = (32.0 * boxes + 48.0 * recs_tested) * scale + H * W * 24
        a_gbs = alg / (kern_ms * 1e-3) / 1e9
        roofline.update({"bound": "hbm", "achieved": a_gbs, "peak": peaks["hbm_gbs"], "unit": "GB/s",
                         "frac": a_gbs / peaks["hbm_gbs"], "peak_source": f"{peak_kind} MEASURED_PEAKS.json hbm_gbs",
                         "issue": {"achieved": achieved / 1e9, "peak": issue_peak / 1e9, "frac": achieved / issue_peak}})
        roofline.pop("hbm", None)
    label = (f"synthetic triangle soup {build.n_tris} tris {W}x{H} {spp}spp depth{MAX_DEPTH}" if is_soup
             else f"scenes/{name}.cu {W}x{H} {spp}spp depth{MAX_DEPTH}")
    out = {"metric": "path Msamples/s", "value": value, "unit": "Msamples/s", "n_gpus": world, "steps": args.steps,
           "warmup": args.warmup, "ms_per_step": ms, "higher_is_better": True, "scaling": "strong", "vs_baseline": None,
           "dtype": "f32", "data": "synthetic",
           "config": {"workload": label, "partition": f"spp/{world}",
                      "l2": "flushed between timed steps (256 MiB write)", "strategy": args.strategy,
                      "prims": build.n_prims, "nodes": build.n_nodes,
                      "chunk_samples": int(st.chunk_samples), "work_items": int(st.work_items)},
           "mrays_per_s": float(rays[0].item()) / ms / 1e3, "mrays_ref_equiv_per_s": float(rays[1].item()) / ms / 1e3,
           "rays_per_sample": float(rays[0].item()) / samples,
           "e2e": {"value": samples / ms_e2e / 1e3, "unit": "Msamples/s", "ms_per_step": ms_e2e, "h2d_bytes_per_step": h2d,
                   "d2h_bytes_per_step": H * W * 12},
           "gpu_launches": int((st.kernel_launches + (1 if world > 1 else 0)) * args.steps),
           "roofline": roofline, "clocks": clocks,
           "build": {"ms": build.build_ms, "max_depth": build.max_depth, "sah": build.sah_cost,
                     "wide_nodes": build.n_wide_nodes, "wide_depth": build.wide_depth}}
    if not args.no_cpu_baseline and world == 1 and is_soup:
        out["cpu_baseline"] = {"unavailable": "the host reference build scans meshes with 2048-triangle leaves (bvh.cuh:71-98): "
                                              "no bounded sample of a 1M+ triangle frame finishes in the bench budget"}
    elif not args.no_cpu_baseline and world == 1:
        out["cpu_baseline"] = cpu_baseline(name, H, W, spp)
        out["ref_cuda_baseline"] = ref_cuda_baseline(name, H, W, scene)
    print(json.dumps(out))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
