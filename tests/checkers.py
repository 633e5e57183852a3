"""ctypes access to the two CPU checkers (test infrastructure):
  oracle/librt_oracle.so       our restatement of the reference path
  oracle/_ref/libref_host.so   the reference's own sources compiled for the host
Both are built by `make -C oracle` (also run by __graft_entry__.build())."""
import ctypes as C
import importlib
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
pkg = importlib.import_module("ray-tracing-cuda_b200")
abi = pkg.abi

# RT_ORACLE_LIB: another build of the same restatement (tools/sanitize_oracle.sh: the ASan + UBSan build)
_ORACLE = os.environ.get("RT_ORACLE_LIB") or os.path.join(ROOT, "oracle", "librt_oracle.so")
_REF = os.path.join(ROOT, "oracle", "_ref", "libref_host.so")


def _build():
    subprocess.run(["make", "-s", "-C", os.path.join(ROOT, "oracle")], check=True, stdout=subprocess.DEVNULL)


def oracle_lib():
    if not os.path.exists(_ORACLE):
        _build()
    lib = C.CDLL(_ORACLE)
    lib.ro_scene_create.restype = C.c_void_p
    lib.ro_scene_create.argtypes = [C.POINTER(abi.RtSceneDesc)]
    lib.ro_scene_destroy.argtypes = [C.c_void_p]
    lib.ro_mesh_sorted_faces.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.ro_camera_look_at.argtypes = [C.c_void_p] * 3 + [C.c_double] * 4 + [C.c_void_p]
    lib.ro_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_int, C.c_void_p, C.c_int]
    lib.ro_primary_rays.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_void_p]
    lib.ro_render_reference.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                        C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_void_p, C.c_int]
    lib.ro_render_product.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_int, C.c_uint64,
                                      C.c_void_p, C.c_void_p, C.c_int]
    lib.ro_render_product_fixed.argtypes = lib.ro_render_product.argtypes
    lib.ro_fixed_to_float.argtypes = [C.c_void_p, C.c_int64, C.c_void_p]
    lib.ro_sphere_direction.argtypes = [C.c_float, C.c_float, C.c_void_p]
    lib.ro_cbrt_unit.restype = C.c_float
    lib.ro_cbrt_unit.argtypes = [C.c_float]
    lib.ro_philox4x32_10.argtypes = [C.c_void_p] * 3
    lib.ro_curand_uniform_seq.argtypes = [C.c_uint64, C.c_uint64, C.c_int, C.c_void_p]
    return lib


def set_pow_mode(libm):
    """1: the oracle calls libm pow() like the host build of the reference; 0: x*x / sqrt."""
    oracle_lib().ro_set_pow_mode(int(libm))


def have_ref():
    if not os.path.exists(_REF) and os.path.isdir("/root/reference/ray-tracing-cuda"):
        _build()
    return os.path.exists(_REF)


def ref_lib():
    lib = C.CDLL(_REF)
    lib.ref_scene_create_named.restype = C.c_void_p
    lib.ref_scene_create_named.argtypes = [C.c_char_p, C.c_uint64, C.c_void_p, C.c_int, C.c_void_p, C.c_int, C.c_int]
    lib.ref_scene_create.restype = C.c_void_p
    lib.ref_scene_create.argtypes = [C.POINTER(abi.RtSceneDesc), C.c_void_p]
    lib.ref_camera_look_at.argtypes = [C.c_void_p] * 3 + [C.c_double] * 4 + [C.c_void_p]
    lib.ref_scene_camera.argtypes = [C.c_void_p, C.c_void_p]
    lib.ref_scene_destroy.argtypes = [C.c_void_p]
    lib.ref_scene_counts.argtypes = [C.c_void_p] + [C.POINTER(C.c_int)] * 4
    lib.ref_scene_describe.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p]
    lib.ref_scene_mesh_info.argtypes = [C.c_void_p, C.c_int, C.POINTER(C.c_int)]
    lib.ref_scene_mesh_faces.argtypes = [C.c_void_p, C.c_int, C.c_void_p]
    lib.ref_trace_closest.restype = C.c_int64
    lib.ref_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_double, C.c_double, C.c_void_p, C.c_int]
    lib.ref_render.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int, C.c_int, C.c_void_p, C.c_int]
    lib.ref_triangle_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_double, C.c_double, C.POINTER(C.c_double), C.c_void_p,
                                     C.POINTER(C.c_double), C.POINTER(C.c_double)]
    lib.ref_aabb_hit.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_double, C.c_double]
    return lib


def camera_from_spec(lib_fn, spec):
    cam = np.zeros((), dtype=abi.camera_dtype)
    lib_fn(spec.position.ctypes.data, spec.look_at.ctypes.data, spec.up.ctypes.data, spec.fov, spec.aspect,
           spec.aperture, spec.focus_distance, cam.ctypes.data)
    return cam


class OracleScene:
    """Our restatement (oracle/rt_oracle.cpp) bound to one SceneData."""

    def __init__(self, scene, camera=None):
        self.lib = oracle_lib()
        self.scene = scene
        d, self._keep = scene.desc()
        self.h = self.lib.ro_scene_create(C.byref(d))
        self.camera = camera

    def camera_look_at(self, spec):
        self.camera = camera_from_spec(self.lib.ro_camera_look_at, spec)
        return self.camera

    def trace(self, rays, t_from=1e-3, t_to=float("inf"), quirks=1, nthreads=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(len(rays), dtype=abi.hit_dtype)
        self.lib.ro_trace_closest(self.h, rays.ctypes.data, len(rays), t_from, t_to, quirks, out.ctypes.data, nthreads)
        return out

    def primary_rays(self, height, width):
        rays = np.zeros((height * width, 6), np.float32)
        self.lib.ro_primary_rays(self.camera.ctypes.data, height, width, rays.ctypes.data)
        return rays

    def sorted_faces(self, mesh=0):
        n = len(self.scene.meshes[mesh])
        out = np.zeros((n, 9), np.float32)
        self.lib.ro_mesh_sorted_faces(self.h, mesh, out.ctypes.data)
        return out

    def render_reference(self, height, width, spp, seed, depth_limit=10, state0_draws=0, quirks=1, post=0, nthreads=0):
        out = np.zeros((height, width, 3), np.float32)
        cnt = np.zeros(3, np.uint64)
        self.lib.ro_render_reference(self.h, self.camera.ctypes.data, height, width, spp, depth_limit, seed,
                                     state0_draws, quirks, post, out.ctypes.data, cnt.ctypes.data, nthreads)
        return out, cnt

    def render_product(self, height, width, spp_begin, spp_count, seed, depth_limit=10, accum=None, nthreads=0):
        if accum is None:
            accum = np.zeros((height, width, 3), np.float32)
        cnt = np.zeros(3, np.uint64)
        self.lib.ro_render_product(self.h, self.camera.ctypes.data, height, width, spp_begin, spp_count, depth_limit,
                                   seed, accum.ctypes.data, cnt.ctypes.data, nthreads)
        return accum, cnt

    def render_product_fixed(self, height, width, spp_begin, spp_count, seed, depth_limit=10, accum=None, nthreads=0):
        """The exact sums (int64, 2^-30 units) rt_render_fixed accumulates."""
        if accum is None:
            accum = np.zeros((height, width, 3), np.int64)
        cnt = np.zeros(3, np.uint64)
        self.lib.ro_render_product_fixed(self.h, self.camera.ctypes.data, height, width, spp_begin, spp_count,
                                         depth_limit, seed, accum.ctypes.data, cnt.ctypes.data, nthreads)
        return accum, cnt

    def __del__(self):
        try:
            self.lib.ro_scene_destroy(self.h)
        except Exception:
            pass


class RefScene:
    """The reference's own code on the host (oracle/_ref/libref_host.so)."""

    def __init__(self, handle, lib):
        self.lib, self.h = lib, handle

    @staticmethod
    def named(name, seed=10086, faces9=None, texture=None):
        lib = ref_lib()
        f = None if faces9 is None else np.ascontiguousarray(faces9, np.float32)
        t = None if texture is None else np.ascontiguousarray(texture, np.uint8)
        h = lib.ref_scene_create_named(name.encode(), seed, None if f is None else f.ctypes.data,
                                       0 if f is None else len(f), None if t is None else t.ctypes.data,
                                       0 if t is None else t.shape[0], 0 if t is None else t.shape[1])
        assert h, name
        return RefScene(h, lib)

    @staticmethod
    def from_scene(scene, camera=None):
        lib = ref_lib()
        d, keep = scene.desc()
        h = lib.ref_scene_create(C.byref(d), None if camera is None else camera.ctypes.data)
        return RefScene(h, lib)

    def camera(self):
        cam = np.zeros((), dtype=abi.camera_dtype)
        self.lib.ref_scene_camera(self.h, cam.ctypes.data)
        return cam

    def describe(self):
        """Flatten the reference object graph into a SceneData."""
        n = [C.c_int() for _ in range(4)]
        self.lib.ref_scene_counts(self.h, *[C.byref(x) for x in n])
        hit = np.zeros(n[0].value, dtype=abi.hitable_dtype)
        mat = np.zeros(n[1].value, dtype=abi.material_dtype)
        self.lib.ref_scene_describe(self.h, hit.ctypes.data, mat.ctypes.data)
        meshes = []
        for i in range(n[2].value):
            uv = C.c_int()
            nf = self.lib.ref_scene_mesh_info(self.h, i, C.byref(uv))
            f = np.zeros((nf, 15 if uv.value else 9), np.float32)
            self.lib.ref_scene_mesh_faces(self.h, i, f.ctypes.data)
            meshes.append(f)
        return abi.SceneData(hit, mat, meshes, camera=self.camera())

    def trace(self, rays, t_from=1e-3, t_to=float("inf"), nthreads=0):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 6)
        out = np.zeros(len(rays), dtype=abi.hit_dtype)
        bad = self.lib.ref_trace_closest(self.h, rays.ctypes.data, len(rays), t_from, t_to, out.ctypes.data, nthreads)
        assert bad == 0, f"id recovery disagreed with world->Hit on {bad} rays"
        return out

    def render(self, height, width, spp, post=0, nthreads=0):
        out = np.zeros((height, width, 3), np.float32)
        self.lib.ref_render(self.h, height, width, spp, post, out.ctypes.data, nthreads)
        return out


def random_rays(n, seed, lo, hi, toward=None, spread=1.0):
    """Seeded rays with origins uniform in [lo,hi]^3 and (glm-normalized, float) directions."""
    rng = np.random.RandomState(seed)
    o = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    if toward is None:
        d = rng.normal(size=(n, 3))
    else:
        d = (np.asarray(toward, np.float64)[None, :] - o) + rng.normal(size=(n, 3)) * spread
    d = d.astype(np.float32)
    # glm normalize: v * (1 / sqrt(dot)) in float
    dot = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32) + (d[:, 2] * d[:, 2]).astype(np.float32)
    inv = (np.float32(1) / np.sqrt(dot.astype(np.float32))).astype(np.float32)
    d = (d * inv[:, None]).astype(np.float32)
    return np.concatenate([o, d], 1).astype(np.float32)


def degenerate_rays(n, seed, lo, hi):
    """Seeded rays that stress the slab tests: one or two direction components exactly +-0 (infinite
    reciprocal), origins on a grid that includes the box planes lo / hi exactly, and tiny components."""
    rng = np.random.RandomState(seed)
    grid = np.linspace(lo, hi, 9).astype(np.float32)
    o = grid[rng.randint(0, 9, (n, 3))].astype(np.float32)
    jitter = rng.uniform(-1, 1, (n, 3)).astype(np.float32) * np.float32((hi - lo) * 0.3)
    o = np.where(rng.rand(n, 3) < 0.5, o, o + jitter).astype(np.float32)
    d = rng.normal(size=(n, 3)).astype(np.float32)
    kind = rng.randint(0, 4, n)
    for i in range(n):
        ax = rng.permutation(3)
        if kind[i] == 0:    # one component exactly 0
            d[i, ax[0]] = 0.0
        elif kind[i] == 1:  # axis parallel, either sign, -0 included
            d[i, ax[0]] = 0.0
            d[i, ax[1]] = -0.0
        elif kind[i] == 2:  # one component far below the others
            d[i, ax[0]] = np.float32(1e-30) * np.float32(rng.choice([-1, 1]))
        else:               # denormal component
            d[i, ax[0]] = np.float32(1e-42)
    dot = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32) + (d[:, 2] * d[:, 2]).astype(np.float32)
    inv = (np.float32(1) / np.sqrt(dot.astype(np.float32))).astype(np.float32)
    d = (d * inv[:, None]).astype(np.float32)
    return np.concatenate([o, d], 1).astype(np.float32)


def hits_equal(a, b, fields=("t", "obj", "sub", "u", "v", "normal", "material")):
    """Bitwise comparison of two rt_hit arrays; returns indices that differ."""
    bad = np.zeros(len(a), bool)
    for f in fields:
        x, y = a[f], b[f]
        if x.dtype.kind == "f":
            x = x.view(np.uint64 if x.dtype.itemsize == 8 else np.uint32)
            y = y.view(np.uint64 if y.dtype.itemsize == 8 else np.uint32)
        d = x != y
        if d.ndim > 1:
            d = d.any(axis=1)
        bad |= d
    return np.nonzero(bad)[0]


def textured_mesh_scene(n_faces=3000, seed=7):
    """A Face<true> mesh (bvh.cuh:35-48: per-corner texcoords, interpolated with the hit's barycentrics) behind a
    Lambertian(ImageTexture) — texcoords in [-0.5, 1.5] so that the wrap rule of image_texture.cu:9-38 is exercised —
    plus the sky.  No shipped scene has one (birthday textures a sphere), so this is the only coverage of that path."""
    scenes, abi = pkg.scenes, pkg.abi
    rng = np.random.RandomState(seed)
    f9 = scenes.soup_mesh(n_faces, seed=5) * np.float32(2.0)
    uv = rng.uniform(-0.5, 1.5, (len(f9), 6)).astype(np.float32)
    tex = rng.randint(0, 256, (16, 32, 4)).astype(np.uint8)
    mats = [scenes._mat(abi.RT_MAT_LAMBERTIAN, (0, 0, 0), texture=0), scenes._mat(abi.RT_MAT_SKY)]
    hs = [scenes._hit(abi.RT_HIT_MESH, 0, mesh=0), scenes._hit(abi.RT_HIT_SKY, 1)]
    scene = abi.SceneData(np.stack(hs), np.stack(mats), meshes=[np.concatenate([f9, uv], 1)], textures=[tex])
    spec = scenes.CameraSpec((0, 0, -6), (0, 0, 0), (0, 1, 0), 0.7, 16 / 9)
    return scene, spec
