"""rt_render_static (include/rtcore.h): rt_render with its scratch in the library's static device array.

No counterpart in the reference.  The entry point must give the bits rt_render gives (both round the same exact
integer sums once), on two calls in a row (the array keeps the previous call's contents: the second call must not
see them), for an accumulating call, and must refuse frames larger than RT_STATIC_MAX_PIXELS.  Collected last on
purpose (the file name sorts after every other test file).
"""
import numpy as np
import pytest

import checkers as ck

pytestmark = pytest.mark.gpu

rt = ck.pkg
scenes = ck.pkg.scenes
RT_STATIC_MAX_PIXELS = 262144


def _bits(t):
    return t.cpu().numpy().view(np.uint32)


@pytest.mark.parametrize("name", ["cornell_box", "spheres"])
def test_static_scratch_render_equals_rt_render(name):
    import torch

    if name == "cornell_box":
        s, c = scenes.cornell_box()
    else:
        s, c, _ = scenes.spheres(10086, "rtl")
    H, W, spp = 60, 80, 5
    ctx = rt.Context(0)
    try:
        ctx.commit(s)
        ctx.set_camera(rt.core.camera_look_at(c))
        want = torch.zeros(H, W, 3, device="cuda:0")
        got = torch.full((H, W, 3), 123.0, device="cuda:0")  # CLEAR must overwrite this
        CLEAR = rt.abi.RT_RENDER_CLEAR
        # two calls in a row with different seeds: the second starts from the first one's leftovers in the array
        for seed in (3, 4):
            ctx.render_device(want.data_ptr(), H, W, spp, 0, 10, seed=seed, flags=CLEAR)
            ctx.render_static_device(got.data_ptr(), H, W, spp, 0, 10, seed=seed, flags=CLEAR)
            torch.cuda.synchronize()
            assert float(want.abs().max()) > 0
            assert (_bits(got) == _bits(want)).all(), f"seed {seed}: static-scratch render differs from rt_render"
        # a smaller frame after a larger one, accumulating (no CLEAR) on top of the caller's buffer, other sample range
        h2, w2 = 24, 40
        want2 = torch.ones(h2, w2, 3, device="cuda:0")
        got2 = torch.ones(h2, w2, 3, device="cuda:0")
        ctx.render_device(want2.data_ptr(), h2, w2, 3, 2, 10, seed=9, flags=0)
        ctx.render_static_device(got2.data_ptr(), h2, w2, 3, 2, 10, seed=9, flags=0)
        torch.cuda.synchronize()
        assert (_bits(got2) == _bits(want2)).all()
        assert float((want2 - 1).abs().max()) > 0
    finally:
        ctx.close()


def test_static_scratch_limits_and_flags():
    import torch

    s, c = scenes.cornell_box()
    ctx = rt.Context(0)
    try:
        ctx.commit(s)
        ctx.set_camera(rt.core.camera_look_at(c))
        buf = torch.zeros(513 * 512 * 3, device="cuda:0")
        # exactly the maximum is accepted
        ctx.render_static_device(buf.data_ptr(), 512, 512, 1, 0, 10, seed=1, flags=rt.abi.RT_RENDER_CLEAR)
        torch.cuda.synchronize()
        assert 512 * 512 == RT_STATIC_MAX_PIXELS and float(buf.abs().max()) > 0
        with pytest.raises(rt.RtError, match="RT_STATIC_MAX_PIXELS"):
            ctx.render_static_device(buf.data_ptr(), 513, 512, 1, 0, 10, seed=1, flags=rt.abi.RT_RENDER_CLEAR)
        with pytest.raises(rt.RtError, match="wavefront"):
            ctx.render_static_device(buf.data_ptr(), 16, 16, 1, 0, 10, seed=1, flags=rt.abi.RT_RENDER_WAVEFRONT)
        # and the array still works afterwards
        want = torch.zeros(16, 16, 3, device="cuda:0")
        got = torch.zeros(16, 16, 3, device="cuda:0")
        ctx.render_device(want.data_ptr(), 16, 16, 2, 0, 10, seed=2, flags=rt.abi.RT_RENDER_CLEAR)
        ctx.render_static_device(got.data_ptr(), 16, 16, 2, 0, 10, seed=2, flags=rt.abi.RT_RENDER_CLEAR)
        torch.cuda.synchronize()
        assert (_bits(got) == _bits(want)).all()
    finally:
        ctx.close()
