"""ctypes binding of librtcore.so (the C-ABI declared in include/rtcore.h).

Host-side mirror of the reference driver for Python callers (tests, bench): `Context` owns an
rt_ctx, `commit` replaces scene construction + BVH build, `render` replaces the PathTracing launch.
Nothing here computes: a missing or unloadable library is a hard error.
"""
import ctypes as C
import os

import numpy as np

from . import abi

_HERE = os.path.dirname(os.path.abspath(__file__))
# RT_CORE_LIB: an experiment build of the same ABI (csrc/Makefile OUT=...); diagnostics only
LIB_PATH = os.environ.get("RT_CORE_LIB") or os.path.join(_HERE, "librtcore.so")
_lib = None


class RtError(RuntimeError):
    pass


def lib():
    """Load librtcore.so (built in-tree by __graft_entry__.build() / `make -C ray-tracing-cuda_b200`)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RtError(
            f"{LIB_PATH} is missing: the CUDA core is not built (run __graft_entry__.build()). "
            "There is no CPU fallback."
        )
    L = C.CDLL(LIB_PATH, mode=C.RTLD_GLOBAL)
    vp, i32, i64, f64 = C.c_void_p, C.c_int32, C.c_int64, C.c_double
    L.rt_abi_version.restype = C.c_int
    L.rt_last_error_string.restype = C.c_char_p
    L.rt_ctx_create.argtypes = [C.c_int, C.POINTER(vp)]
    L.rt_ctx_destroy.argtypes = [vp]
    L.rt_scene_commit.argtypes = [vp, C.POINTER(abi.RtSceneDesc), C.POINTER(abi.RtBuildStats)]
    L.rt_camera_set.argtypes = [vp, vp]
    L.rt_camera_look_at.argtypes = [vp, vp, vp, f64, f64, f64, f64, vp]
    L.rt_trace_closest_host.argtypes = [vp, vp, i64, f64, f64, C.c_int, vp]
    L.rt_trace_closest_device.argtypes = [vp, vp, i64, f64, f64, C.c_int, vp, vp]
    L.rt_primary_rays_host.argtypes = [vp, C.c_int, C.c_int, vp]
    L.rt_render.argtypes = [vp, C.POINTER(abi.RtRenderParams), vp, C.POINTER(abi.RtRenderStats), vp]
    L.rt_render_fixed.argtypes = [vp, C.POINTER(abi.RtRenderParams), vp, C.POINTER(abi.RtRenderStats), vp]
    L.rt_render_static.argtypes = [vp, C.POINTER(abi.RtRenderParams), vp, vp]
    L.rt_resolve.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp]
    L.rt_resolve_fixed.argtypes = [vp, vp, C.c_int, C.c_int, C.c_int, vp, vp, vp, vp]
    L.rt_reduce_fixed.argtypes = [vp, vp, vp, i64, C.c_int, vp]
    L.rt_render_host.argtypes = [vp, C.POINTER(abi.RtRenderParams), vp, vp, C.POINTER(abi.RtRenderStats)]
    L.rt_get_workload.argtypes = [C.c_int, C.c_int, C.c_int, C.POINTER(C.c_int)]
    L.rt_work_plan.argtypes = [C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, vp, C.POINTER(C.c_int32)]
    L.rt_reduce.argtypes = [vp, vp, vp, i64, C.c_int, vp]
    L.rt_peer_frame_create.argtypes = [vp, C.c_int, C.c_int, vp]
    L.rt_peer_frame_open.argtypes = [vp, C.c_int, C.c_int, vp]
    L.rt_peer_frame_buffers.argtypes = [vp, C.POINTER(vp), C.POINTER(vp), C.POINTER(vp)]
    L.rt_reduce_resolve_peers.argtypes = [vp, C.c_int, C.c_int, vp]
    L.rt_peer_frame_check.argtypes = [vp, vp]
    L.rt_peer_frame_destroy.argtypes = [vp]
    L.rt_nccl_unique_id.argtypes = [vp]
    L.rt_nccl_init.argtypes = [vp, vp, C.c_int, C.c_int]
    L.rt_nccl_destroy.argtypes = [vp]
    L.rt_write_jpeg.argtypes = [C.c_char_p, C.c_int, C.c_int, vp, C.c_int]
    L.rt_write_ppm.argtypes = [C.c_char_p, C.c_int, C.c_int, vp]
    L.rt_write_pfm.argtypes = [C.c_char_p, C.c_int, C.c_int, vp]
    if L.rt_abi_version() != 4:
        raise RtError("librtcore.so ABI version mismatch")
    _lib = L
    return L


def _check(rc):
    if rc != 0:
        raise RtError(f"rtcore error {rc}: {lib().rt_last_error_string().decode(errors='replace')}")


def camera_look_at(spec):
    """Camera constructors (camera.cu:6-38) -> rt_camera record."""
    cam = np.zeros((), dtype=abi.camera_dtype)
    _check(lib().rt_camera_look_at(spec.position.ctypes.data, spec.look_at.ctypes.data, spec.up.ctypes.data,
                                   spec.fov, spec.aspect, spec.aperture, spec.focus_distance, cam.ctypes.data))
    return cam


def get_workload(rank, world_size, spp):
    """GetWorkload (utils.cu:111-113): (samples of this rank, first global sample index)."""
    b = C.c_int()
    n = lib().rt_get_workload(rank, world_size, spp, C.byref(b))
    return n, b.value


def work_plan(n_sm, blocks_per_sm, height, width, spp_begin, spp_count, bounded_cost):
    """rt_work_plan: (phases [(first item, first sample, samples per chunk, chunks per item)], items of the launch, tiles)."""
    ph = np.zeros((5, 4), np.int32)
    nt = C.c_int32()
    n = lib().rt_work_plan(n_sm, blocks_per_sm, height, width, spp_begin, spp_count, int(bounded_cost), ph.ctypes.data,
                           C.byref(nt))
    if n < 0:
        _check(n)
    assert tuple(ph[n, 1:]) == (spp_begin + spp_count, 1, 1)
    return [tuple(int(x) for x in ph[i]) for i in range(n)], int(ph[n, 0]), nt.value


class Context:
    """One rt_ctx on one CUDA device."""

    def __init__(self, device=0):
        self._h = C.c_void_p()
        _check(lib().rt_ctx_create(device, C.byref(self._h)))
        self.build_stats = None

    def close(self):
        if self._h:
            lib().rt_ctx_destroy(self._h)
            self._h = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def commit(self, scene):
        d, keep = scene.desc()
        st = abi.RtBuildStats()
        _check(lib().rt_scene_commit(self._h, C.byref(d), C.byref(st)))
        self.build_stats = st
        return st

    def set_camera(self, cam):
        cam = np.ascontiguousarray(cam, dtype=abi.camera_dtype)
        _check(lib().rt_camera_set(self._h, cam.ctypes.data))

    def trace_closest(self, rays, t_from=1e-3, t_to=float("inf"), mode=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(len(rays), dtype=abi.hit_dtype)
        _check(lib().rt_trace_closest_host(self._h, rays.ctypes.data, len(rays), t_from, t_to, mode, out.ctypes.data))
        return out

    def primary_rays(self, height, width):
        rays = np.zeros((height * width, 6), np.float32)
        _check(lib().rt_primary_rays_host(self._h, height, width, rays.ctypes.data))
        return rays

    def params(self, height, width, spp_count, spp_begin=0, max_depth=10, seed=0, flags=0):
        return abi.RtRenderParams(height, width, spp_begin, spp_count, max_depth, flags, seed)

    def render_host(self, height, width, spp_count, spp_begin=0, max_depth=10, seed=0, flags=0, want_rgb=False):
        """End-to-end through host buffers: returns (accum float32 [H,W,3], rgb or None, stats)."""
        p = self.params(height, width, spp_count, spp_begin, max_depth, seed, flags)
        accum = np.zeros((height, width, 3), np.float32)
        rgb = np.zeros((height, width, 3), np.float32) if want_rgb else None
        st = abi.RtRenderStats()
        _check(lib().rt_render_host(self._h, C.byref(p), accum.ctypes.data, None if rgb is None else rgb.ctypes.data,
                                    C.byref(st)))
        return accum, rgb, st

    def render_device(self, d_accum_ptr, height, width, spp_count, spp_begin=0, max_depth=10, seed=0, flags=0,
                      stream=None):
        """rt_render into a device accumulation buffer (raw pointer, e.g. torch tensor .data_ptr())."""
        p = self.params(height, width, spp_count, spp_begin, max_depth, seed, flags)
        st = abi.RtRenderStats()
        _check(lib().rt_render(self._h, C.byref(p), d_accum_ptr, C.byref(st), stream))
        return st

    def render_fixed_device(self, d_fixed_ptr, height, width, spp_count, spp_begin=0, max_depth=10, seed=0, flags=0,
                            stream=None):
        """rt_render_fixed into caller-owned exact sums (int64 [H, W, 3] on the device, units of 2^-30)."""
        p = self.params(height, width, spp_count, spp_begin, max_depth, seed, flags)
        st = abi.RtRenderStats()
        _check(lib().rt_render_fixed(self._h, C.byref(p), d_fixed_ptr, C.byref(st), stream))
        return st

    def render_static_device(self, d_accum_ptr, height, width, spp_count, spp_begin=0, max_depth=10, seed=0, flags=0,
                             stream=None):
        """rt_render_static: rt_render with its scratch in the library's static device array (no allocation)."""
        p = self.params(height, width, spp_count, spp_begin, max_depth, seed, flags)
        _check(lib().rt_render_static(self._h, C.byref(p), d_accum_ptr, stream))

    def resolve_fixed_device(self, d_fixed_ptr, height, width, spp_total, d_accum_ptr=None, d_rgb_ptr=None,
                             d_rgb8_ptr=None, stream=None):
        _check(lib().rt_resolve_fixed(self._h, d_fixed_ptr, height, width, spp_total, d_accum_ptr, d_rgb_ptr, d_rgb8_ptr,
                                      stream))

    def reduce_fixed(self, nccl_comm, d_fixed_ptr, n_values, root=0, stream=None):
        _check(lib().rt_reduce_fixed(self._h, nccl_comm, d_fixed_ptr, n_values, root, stream))

    def resolve_device(self, d_accum_ptr, height, width, spp_total, d_rgb_ptr=None, d_rgb8_ptr=None, stream=None):
        _check(lib().rt_resolve(self._h, d_accum_ptr, height, width, spp_total, d_rgb_ptr, d_rgb8_ptr, stream))

    def reduce(self, nccl_comm, d_accum_ptr, n_floats, root=0, stream=None):
        _check(lib().rt_reduce(self._h, nccl_comm, d_accum_ptr, n_floats, root, stream))

    # ---- the reduce over NVLink peer memory (rt_peer.cu; replaces utils.cu:115-130, 232-241 without NCCL) ----
    def peer_frame_create(self, height, width):
        """Allocate this rank's frame; returns the 64-byte CUDA IPC handle the other ranks open."""
        buf = C.create_string_buffer(64)
        _check(lib().rt_peer_frame_create(self._h, height, width, buf))
        return buf.raw

    def peer_frame_open(self, world_size, rank, handles):
        """handles: the 64-byte handle of every rank, in rank order (own entry is not opened)."""
        blob = C.create_string_buffer(b"".join(bytes(h) for h in handles), 64 * world_size) if world_size > 1 else None
        _check(lib().rt_peer_frame_open(self._h, world_size, rank, blob))

    def peer_frame_buffers(self):
        """Device addresses (exact sums int64, float sums, resolved image) of THIS rank's frame."""
        a, b, c = C.c_void_p(), C.c_void_p(), C.c_void_p()
        _check(lib().rt_peer_frame_buffers(self._h, C.byref(a), C.byref(b), C.byref(c)))
        return a.value, b.value, c.value

    def reduce_resolve_peers(self, spp_total, root=0, stream=None):
        _check(lib().rt_reduce_resolve_peers(self._h, spp_total, root, stream))

    def peer_frame_check(self, stream=None):
        _check(lib().rt_peer_frame_check(self._h, stream))

    def peer_frame_destroy(self):
        lib().rt_peer_frame_destroy(self._h)

    def nccl_init(self, unique_id, world_size, rank):
        """Bind an NCCL communicator to this context from a 128-byte ncclUniqueId (see nccl_unique_id)."""
        buf = C.create_string_buffer(bytes(unique_id), 128)
        _check(lib().rt_nccl_init(self._h, buf, world_size, rank))

    def nccl_destroy(self):
        lib().rt_nccl_destroy(self._h)


def nccl_unique_id():
    buf = C.create_string_buffer(128)
    _check(lib().rt_nccl_unique_id(buf))
    return buf.raw


def write_image(path, rgb8):
    """WriteImage (utils.cu:87-96): rgb8 uint8 [H, W, 3] -> .jpeg / .jpg / .ppm by extension."""
    rgb8 = np.ascontiguousarray(rgb8, np.uint8)
    h, w = rgb8.shape[:2]
    if path.endswith(".ppm"):
        _check(lib().rt_write_ppm(path.encode(), w, h, rgb8.ctypes.data))
    else:
        _check(lib().rt_write_jpeg(path.encode(), w, h, rgb8.ctypes.data, 100))


def write_pfm(path, rgb):
    rgb = np.ascontiguousarray(rgb, np.float32)
    h, w = rgb.shape[:2]
    _check(lib().rt_write_pfm(path.encode(), w, h, rgb.ctypes.data))
