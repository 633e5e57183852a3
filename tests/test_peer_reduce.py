"""The reduce over NVLink peer frames (csrc/rt_peer.cu; replaces GatherImageData, utils.cu:115-130,232-241).

One GPU: the world_size-1 frame (flags, epochs, resolve arithmetic) against rt_resolve_fixed.
Two GPUs: two processes (torch.distributed.run), each renders its GetWorkload share into its frame; the root's frame
must hold the image ONE rank renders, byte for byte, over several frames (tests/peer_worker.py).
"""
import os
import subprocess
import sys

import numpy as np
import pytest

import checkers as ck

rt = ck.pkg
scenes = ck.pkg.scenes
HERE = os.path.dirname(os.path.abspath(__file__))


class _Dev:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


@pytest.mark.gpu
def test_single_rank_frame_equals_plain_resolve():
    import torch

    H, W, spp = 48, 64, 6
    ctx = rt.Context(0)
    try:
        s, c = scenes.cornell_box()
        ctx.commit(s)
        ctx.set_camera(rt.core.camera_look_at(c))
        handle = ctx.peer_frame_create(H, W)
        assert len(handle) == 64 and any(handle)
        ctx.peer_frame_open(1, 0, [handle])
        p_sums, p_accum, p_rgb = ctx.peer_frame_buffers()
        sums = torch.as_tensor(_Dev(p_sums, (H, W, 3), "<i8"), device="cuda:0")
        accum = torch.as_tensor(_Dev(p_accum, (H, W, 3), "<f4"), device="cuda:0")
        rgb = torch.as_tensor(_Dev(p_rgb, (H, W, 3), "<f4"), device="cuda:0")
        ref_sums = torch.zeros(H, W, 3, dtype=torch.int64, device="cuda:0")
        ref_accum = torch.zeros(H, W, 3, device="cuda:0")
        ref_rgb = torch.zeros(H, W, 3, device="cuda:0")
        for frame in range(3):  # three epochs of the ready / done counters
            seed = 5 + frame
            ctx.render_fixed_device(p_sums, H, W, spp, 0, 10, seed=seed, flags=rt.abi.RT_RENDER_CLEAR)
            ctx.reduce_resolve_peers(spp, 0)
            ctx.peer_frame_check()
            ctx.render_fixed_device(ref_sums.data_ptr(), H, W, spp, 0, 10, seed=seed, flags=rt.abi.RT_RENDER_CLEAR)
            ctx.resolve_fixed_device(ref_sums.data_ptr(), H, W, spp, ref_accum.data_ptr(), ref_rgb.data_ptr())
            torch.cuda.synchronize()
            assert torch.equal(sums, ref_sums)
            assert accum.cpu().numpy().tobytes() == ref_accum.cpu().numpy().tobytes()
            assert rgb.cpu().numpy().tobytes() == ref_rgb.cpu().numpy().tobytes()
            assert float(ref_rgb.max()) > 0
        del sums, accum, rgb
        ctx.peer_frame_destroy()
    finally:
        ctx.close()


@pytest.mark.gpu
def test_peer_frame_call_order_errors():
    ctx = rt.Context(0)
    try:
        with pytest.raises(rt.RtError):
            ctx.peer_frame_buffers()  # no frame yet
        with pytest.raises(rt.RtError):
            ctx.reduce_resolve_peers(4, 0)
        ctx.peer_frame_create(8, 8)
        with pytest.raises(rt.RtError):
            ctx.reduce_resolve_peers(4, 0)  # not opened
        with pytest.raises(rt.RtError):
            ctx.peer_frame_open(17, 0, [b"\0" * 64] * 17)  # more than 16 ranks
        ctx.peer_frame_open(1, 0, None)
        with pytest.raises(rt.RtError):
            ctx.reduce_resolve_peers(4, 1)  # root outside the world
    finally:
        ctx.close()


@pytest.mark.gpu
def test_two_ranks_over_peer_frames_render_the_one_rank_image(tmp_path):
    import torch

    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, MASTER_ADDR="127.0.0.1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr", "127.0.0.1",
           "--master-port", "29533", os.path.join(HERE, "peer_worker.py"), str(tmp_path / "out.npz")]
    r = subprocess.run(cmd, env=env, capture_output=True, text=True, timeout=600)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    d = np.load(tmp_path / "out.npz")
    assert int(d["frames"]) == 3
    assert int(d["mismatching_values"]) == 0
    assert float(d["max_rgb"]) > 0
