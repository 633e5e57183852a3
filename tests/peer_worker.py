"""One rank of tests/test_peer_reduce.py::test_two_ranks_over_peer_frames_render_the_one_rank_image.

Launched by torch.distributed.run with two processes.  Every rank renders GetWorkload(rank, world, spp) samples of a
small Cornell box into its peer frame and calls rt_reduce_resolve_peers; rank 0 then renders ALL samples alone and
compares its frame (exact sums, float sums, resolved image) with that, for three frames with different seeds.
"""
import importlib
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


class _Dev:
    def __init__(self, ptr, shape, typestr):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False), "version": 2}


def main(out_path):
    import torch
    import torch.distributed as dist

    rt = importlib.import_module("ray-tracing-cuda_b200")
    rdist = importlib.import_module("ray-tracing-cuda_b200.distributed")
    rank, world, local = rdist.init_process_group()
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    H, W, spp = 96, 128, 9  # 9 samples over 2 ranks: 5 + 4
    ctx = rt.Context(local)
    s, c = rt.scenes.cornell_box()
    ctx.commit(s)
    ctx.set_camera(rt.core.camera_look_at(c))
    p_sums, p_accum, p_rgb = rdist.open_peer_frame(ctx, H, W)
    count, begin = rdist.partition(spp, rank, world)
    mismatches, max_rgb = 0, 0.0
    for frame in range(3):
        seed = 11 + frame
        ctx.render_fixed_device(p_sums, H, W, count, begin, 10, seed=seed, flags=rt.abi.RT_RENDER_CLEAR)
        ctx.reduce_resolve_peers(spp, 0)
        ctx.peer_frame_check()
        if rank == 0:
            got = [torch.as_tensor(_Dev(p, (H, W, 3), t), device=dev).clone()
                   for p, t in ((p_sums, "<i8"), (p_accum, "<f4"), (p_rgb, "<f4"))]
            ref_sums = torch.zeros(H, W, 3, dtype=torch.int64, device=dev)
            ref_accum = torch.zeros(H, W, 3, device=dev)
            ref_rgb = torch.zeros(H, W, 3, device=dev)
            ctx.render_fixed_device(ref_sums.data_ptr(), H, W, spp, 0, 10, seed=seed, flags=rt.abi.RT_RENDER_CLEAR)
            ctx.resolve_fixed_device(ref_sums.data_ptr(), H, W, spp, ref_accum.data_ptr(), ref_rgb.data_ptr())
            torch.cuda.synchronize()
            mismatches += int((got[0] != ref_sums).sum())
            mismatches += int((got[1].view(torch.int32) != ref_accum.view(torch.int32)).sum())
            mismatches += int((got[2].view(torch.int32) != ref_rgb.view(torch.int32)).sum())
            max_rgb = max(max_rgb, float(ref_rgb.max()))
        dist.barrier()
    ctx.peer_frame_destroy()
    ctx.close()
    if rank == 0:
        np.savez(out_path, frames=3, mismatching_values=mismatches, max_rgb=max_rgb)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main(sys.argv[1])
