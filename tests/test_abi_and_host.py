"""CPU-only checks of the boundary and the host logic: the C-ABI library loads and exports every symbol that
include/rtcore.h declares, compute entry points fail loudly without a GPU, record layouts match the header, the
Philox generator matches the Random123 known-answer vectors, and the sample partition + reduce of the multi-GPU
path works across two gloo ranks."""
import ctypes as C
import os
import re
import subprocess
import sys

import numpy as np
import pytest

import checkers as ck

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
rt = ck.pkg
abi = rt.abi


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "rtcore.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(rt_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    lib = rt.lib()
    names = _declared_symbols()
    assert len(names) >= 20, names
    for n in names:
        assert hasattr(lib, n), f"librtcore.so lacks {n} declared in include/rtcore.h"
    assert lib.rt_abi_version() == 4


def test_record_sizes_match_header():
    assert abi.hitable_dtype.itemsize == 72 and abi.material_dtype.itemsize == 32
    assert abi.hit_dtype.itemsize == 40 and abi.camera_dtype.itemsize == 96


def test_no_cpu_fallback():
    import torch
    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(rt.RtError, match="no usable CUDA device|CUDA"):
        rt.Context(0)


def test_get_workload_matches_reference_formula():
    for spp in (0, 1, 7, 16, 100, 1024):
        for ws in (1, 2, 3, 4, 8):
            parts = [rt.core.get_workload(r, ws, spp) for r in range(ws)]
            assert [n for n, _ in parts] == [spp // ws + (1 if r < spp % ws else 0) for r in range(ws)]  # utils.cu:111-113
            begins = [b for _, b in parts]
            assert begins == list(np.concatenate([[0], np.cumsum([n for n, _ in parts])[:-1]]))
            if ck.have_ref():
                assert [n for n, _ in parts] == [ck.ref_lib().ref_get_workload(r, ws, spp) for r in range(ws)]


def _samples_of_plan(phases, n_items, n_tiles, spp_begin, spp_end):
    """What hand_out (rt_render.cu) makes of a plan, for ONE pixel of one tile: the sample ranges its sub-units cover.
    Item it of phase p: group = (it - first item) // tiles, sample k0 = first sample + group * r * ch; chunk c of the
    item covers [k0 + c * ch, min(k0 + (c + 1) * ch, end of the phase)) and is dropped when it starts at or beyond that
    end."""
    cover = []
    bounds = [p[0] for p in phases] + [n_items]
    ends = [p[1] for p in phases[1:]] + [spp_end]
    for p, (first_item, k_begin, ch, r) in enumerate(phases):
        n_it = bounds[p + 1] - first_item
        assert n_it % n_tiles == 0
        for group in range(n_it // n_tiles):
            k0 = k_begin + group * r * ch
            for c in range(r):
                kk = k0 + c * ch
                if kk < ends[p]:
                    cover.append((kk, min(kk + ch, ends[p])))
    return cover


def test_work_plan_covers_every_sample_once(monkeypatch):
    """The work plan of a path-kernel launch (rt_work_plan: the host arithmetic behind rt_render's phases of
    (chunk, chunks per item)) hands out every sample of the range exactly once, for the BASELINE frames, one GPU's share
    of an 8-GPU run, ragged sample counts, a non-zero first sample, both scene classes, and explicit RT_PLAN /
    RT_CHUNK lists (ragged last items inside a phase, lists that stop short or overshoot)."""
    for v in ("RT_PLAN", "RT_CHUNK", "RT_ITEM_CHUNKS"):
        monkeypatch.delenv(v, raising=False)
    cases = [(1024, 1024, 0, 1024, 1), (1024, 1024, 896, 128, 1), (1080, 1920, 0, 512, 1), (1080, 1920, 64, 64, 1),
             (225, 400, 0, 16, 0), (1080, 1920, 0, 256, 0), (2160, 3840, 0, 64, 0), (67, 125, 3, 37, 1), (128, 128, 0, 4096, 1),
             (1024, 1024, 0, 1000, 1), (1024, 1024, 0, 63, 1), (33, 65, 0, 1, 1), (1024, 1024, 5, 777, 1)]
    envs = [{}, {"RT_PLAN": "4x2:6,3x3:1,2x1"}, {"RT_PLAN": "16x4"}, {"RT_PLAN": "2x3:4,1x1:1,5x2:2,3x1"},
            {"RT_PLAN": "1x64:3,4x1"}, {"RT_PLAN": "8x1:100000"}, {"RT_PLAN": "garbage"}, {"RT_CHUNK": "3", "RT_ITEM_CHUNKS": "5"},
            {"RT_CHUNK": "16"}]
    seen_cascade = False
    for env in envs:
        for k, v in env.items():
            monkeypatch.setenv(k, v)
        for (H, W, b, n, bounded) in (cases if not env else cases[:2] + cases[7:8] + cases[11:]):
            for minb in (6, 8):
                phases, n_items, n_tiles = rt.core.work_plan(148, minb, H, W, b, n, bounded)
                assert n_tiles == ((W + 7) // 8) * ((H + 3) // 4)
                assert 1 <= len(phases) <= 4 and phases[0][0] == 0 and phases[0][1] == b
                assert all(p[2] >= 1 and p[3] >= 1 for p in phases)
                cover = sorted(_samples_of_plan(phases, n_items, n_tiles, b, b + n))
                assert cover[0][0] == b and cover[-1][1] == b + n, (env, H, W, b, n, phases)
                assert all(cover[i][1] == cover[i + 1][0] for i in range(len(cover) - 1)), (env, H, W, b, n, phases)
                if not env and bounded and n >= 64 and H * W * n // (148 * minb * 128) >= 512:
                    # the geometric cascade of the box-list scenes: item sizes decrease from phase to phase
                    sizes = [p[2] * p[3] for p in phases]
                    assert sizes == sorted(sizes, reverse=True) and len(phases) >= 2 and sizes[-1] <= 4 and sizes[0] >= 16
                    seen_cascade = True
                if not env and not bounded:
                    assert len(phases) == 1 and phases[0][3] == 1 and phases[0][2] <= 4
        for k in env:
            monkeypatch.delenv(k)
    assert seen_cascade
    # Cornell at 1024 spp and at one GPU's share of an 8-GPU run: the plans DESIGN.md quotes
    assert [(p[2], p[3]) for p in rt.core.work_plan(148, 6, 1024, 1024, 0, 1024, 1)[0]] == [(16, 32), (16, 8), (16, 1), (4, 1)]
    assert [(p[2], p[3]) for p in rt.core.work_plan(148, 6, 1024, 1024, 0, 128, 1)[0]] == [(16, 6), (16, 1), (4, 1)]
    with pytest.raises(Exception):
        rt.core.work_plan(148, 6, 0, 10, 0, 1, 1)


def test_philox4x32_10_known_answers():
    """Random123 kat_vectors for philox4x32-10."""
    lib = ck.oracle_lib()
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        c, k, o = np.array(ctr, np.uint32), np.array(key, np.uint32), np.zeros(4, np.uint32)
        lib.ro_philox4x32_10(c.ctypes.data, k.ctypes.data, o.ctypes.data)
        assert tuple(int(x) for x in o) == want


def test_xorwow_restatement_matches_curand_header():
    lib = ck.oracle_lib()
    out = np.zeros(64, np.float32)
    lib.ro_curand_uniform_seq(10086, 0, 64, out.ctypes.data)
    x = rt.scenes.Xorwow(10086)
    assert [float(x.uniform()) for _ in range(64)] == [float(v) for v in out]


_WORKER = r"""
import importlib, os, sys
sys.path.insert(0, {root!r}); sys.path.insert(0, os.path.join({root!r}, "tests"))
import numpy as np, torch.distributed as dist
import checkers as ck
rt = ck.pkg
rd = importlib.import_module("ray-tracing-cuda_b200.distributed")
rank, world, _ = rd.init_process_group("gloo")
scene, cam = rt.scenes.cornell_box()
o = ck.OracleScene(scene); o.camera_look_at(cam)
H, W, spp = 12, 20, 5
count, begin = rd.partition(spp, rank, world)
local, _ = o.render_product(H, W, begin, count, 42)           # the oracle stands in for rt_render on CPU
total = rd.reduce_on_cpu(local, 0)
payload = rd.broadcast_bytes(bytes(range(128)) if rank == 0 else b"", 128, 0)
assert payload == bytes(range(128))
if rank == 0:
    whole, _ = o.render_product(H, W, 0, spp, 42)
    assert (count, begin) == (3, 0)
    assert np.allclose(total, whole, rtol=1e-6, atol=1e-6), np.abs(total - whole).max()
    print("GLOO_OK")
dist.destroy_process_group()
"""


def test_sample_partition_and_reduce_two_gloo_ranks(tmp_path):
    script = tmp_path / "worker.py"
    script.write_text(_WORKER.format(root=ROOT))
    procs = []
    for r in range(2):
        env = dict(os.environ, RANK=str(r), WORLD_SIZE="2", LOCAL_RANK=str(r), MASTER_ADDR="127.0.0.1", MASTER_PORT="29577")
        procs.append(subprocess.Popen([sys.executable, str(script)], env=env, stdout=subprocess.PIPE, stderr=subprocess.PIPE, text=True))
    outs = [p.communicate(timeout=240) for p in procs]
    assert all(p.returncode == 0 for p in procs), outs
    assert "GLOO_OK" in outs[0][0]


def test_image_writers_round_trip(tmp_path):
    """rt_write_ppm / rt_write_pfm / rt_write_jpeg (the WriteImage replacement, utils.cu:87-96): PPM and PFM
    read back exactly, the baseline JPEG at quality 100 decodes (PIL) to within a few levels of the input,
    and a bad path is an error code, not an abort."""
    rng = np.random.default_rng(5)
    h, w = 37, 53  # not multiples of the 8x8 JPEG block: exercises edge padding
    yy, xx = np.mgrid[0:h, 0:w]
    rgb8 = np.stack([(xx * 255 // (w - 1)), (yy * 255 // (h - 1)), ((xx + yy) * 255 // (h + w - 2))], -1).astype(np.uint8)

    p = str(tmp_path / "a.ppm")
    rt.write_image(p, rgb8)
    raw = open(p, "rb").read()
    head = f"P6\n{w} {h}\n255\n".encode()
    assert raw.startswith(b"P6") and len(raw) >= w * h * 3
    assert np.array_equal(np.frombuffer(raw[-w * h * 3:], np.uint8).reshape(h, w, 3), rgb8)
    assert raw[: len(raw) - w * h * 3].split() == head.split()

    f = rng.standard_normal((h, w, 3)).astype(np.float32)
    p = str(tmp_path / "a.pfm")
    rt.write_pfm(p, f)
    raw = open(p, "rb").read()
    toks = raw[: len(raw) - w * h * 12].split()
    assert toks[0] == b"PF" and [int(toks[1]), int(toks[2])] == [w, h] and float(toks[3]) < 0  # little endian
    back = np.frombuffer(raw[-w * h * 12:], "<f4").reshape(h, w, 3)[::-1]  # PFM rows are bottom-to-top
    assert np.array_equal(back.view(np.uint32), f.view(np.uint32))

    PIL = pytest.importorskip("PIL.Image")
    p = str(tmp_path / "a.jpeg")
    rt.write_image(p, rgb8)
    dec = np.asarray(PIL.open(p).convert("RGB"))
    assert dec.shape == rgb8.shape
    err = np.abs(dec.astype(np.int32) - rgb8.astype(np.int32))
    assert err.mean() < 2.0 and err.max() <= 12, (err.mean(), err.max())

    with pytest.raises(rt.RtError):
        rt.write_image(str(tmp_path / "no_such_dir" / "a.ppm"), rgb8)
    with pytest.raises(rt.RtError):
        rt.write_pfm(str(tmp_path / "no_such_dir" / "a.pfm"), f)
