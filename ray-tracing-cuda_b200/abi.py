"""Binary layouts of include/rtcore.h as numpy dtypes / ctypes structures.

Only layouts live here (no compute): the same records are handed to librtcore.so (product) and,
in tests, to the CPU checkers under oracle/.
"""
import ctypes as C

import numpy as np

RT_MAT_LAMBERTIAN, RT_MAT_METAL, RT_MAT_DIELECTRIC, RT_MAT_DIFFUSE_LIGHT, RT_MAT_SKY = range(5)
RT_HIT_SPHERE, RT_HIT_TRIANGLE, RT_HIT_PARALLELOGRAM, RT_HIT_SKY, RT_HIT_MESH = range(5)

RT_RENDER_CLEAR = 1
RT_RENDER_PROFILE = 2
RT_RENDER_FUSED = 4
RT_RENDER_SORT = 8
RT_RENDER_WAVEFRONT = 16
RT_RENDER_RAYSORT = 128
# rt_trace_closest modes: the scene's render traversal, HitableList::Hit as a sequential scan, and each traversal by name
TRACE_DEFAULT, TRACE_SEQUENTIAL, TRACE_IFIF, TRACE_WHILE_WHILE, TRACE_BOX_LIST = range(5)
FIXED_ONE = 1 << 30  # units of the exact radiance sums (rt_render_fixed)

material_dtype = np.dtype(
    {
        "names": ["type", "texture", "color", "fuzz", "ior"],
        "formats": ["<i4", "<i4", ("<f4", 3), "<f4", "<f8"],
        "offsets": [0, 4, 8, 20, 24],
        "itemsize": 32,
    }
)
hitable_dtype = np.dtype(
    {
        "names": ["type", "material", "p", "radius", "mesh", "group"],
        "formats": ["<i4", "<i4", ("<f4", (4, 3)), "<f8", "<i4", "<i4"],
        "offsets": [0, 4, 8, 56, 64, 68],
        "itemsize": 72,
    }
)
hit_dtype = np.dtype(
    {
        "names": ["t", "obj", "sub", "u", "v", "normal", "material"],
        "formats": ["<f8", "<i4", "<i4", "<f4", "<f4", ("<f4", 3), "<i4"],
        "offsets": [0, 8, 12, 16, 20, 24, 36],
        "itemsize": 40,
    }
)
camera_dtype = np.dtype(
    {
        "names": ["position", "lower_left_corner", "horizontal", "vertical", "u", "v", "w", "is_defocus", "lens_radius"],
        "formats": [("<f4", 3)] * 7 + ["<i4", "<f8"],
        "offsets": [0, 12, 24, 36, 48, 60, 72, 84, 88],
        "itemsize": 96,
    }
)


class RtMesh(C.Structure):
    _fields_ = [("h_faces", C.c_void_p), ("n_faces", C.c_int32), ("has_texcoord", C.c_int32)]


class RtTexture(C.Structure):
    _fields_ = [("h_rgba", C.c_void_p), ("height", C.c_int32), ("width", C.c_int32), ("pitch_bytes", C.c_int64)]


class RtSceneDesc(C.Structure):
    _fields_ = [
        ("hitables", C.c_void_p),
        ("n_hitables", C.c_int32),
        ("materials", C.c_void_p),
        ("n_materials", C.c_int32),
        ("meshes", C.POINTER(RtMesh)),
        ("n_meshes", C.c_int32),
        ("textures", C.POINTER(RtTexture)),
        ("n_textures", C.c_int32),
    ]


class RtBuildStats(C.Structure):
    _fields_ = [
        ("n_prims", C.c_int32),
        ("n_nodes", C.c_int32),
        ("n_spheres", C.c_int32),
        ("n_quads", C.c_int32),
        ("n_tris", C.c_int32),
        ("max_depth", C.c_int32),
        ("build_ms", C.c_float),
        ("upload_ms", C.c_float),
        ("sah_cost", C.c_float),
        ("reserved0_", C.c_int32),
        ("reserved1_", C.c_int32),
    ]


class RtRenderStats(C.Structure):
    _fields_ = [
        ("samples", C.c_uint64),
        ("rays", C.c_uint64),
        ("rays_ref_equiv", C.c_uint64),
        ("node_visits", C.c_uint64),
        ("prim_tests", C.c_uint64),
        ("iterations", C.c_uint32),
        ("kernel_launches", C.c_uint32),
        ("render_ms", C.c_float),
        ("chunk_samples", C.c_uint32),
        ("work_items", C.c_uint32),
        ("seq_fallbacks", C.c_uint32),
        ("max_box_tests_per_ray", C.c_uint32),
        ("reserved_", C.c_uint32),
    ]

    def as_dict(self):
        return {k: getattr(self, k) for k, _ in self._fields_}


class RtRenderParams(C.Structure):
    _fields_ = [
        ("height", C.c_int32),
        ("width", C.c_int32),
        ("spp_begin", C.c_int32),
        ("spp_count", C.c_int32),
        ("max_depth", C.c_int32),
        ("flags", C.c_int32),
        ("seed", C.c_uint64),
    ]


class SceneData:
    """A scene as plain arrays: the host-side mirror of what the reference builds with device-side
    `new` + HitableList::Append.  meshes: list of (faces float32 [n, 9|15]); textures: list of
    uint8 [h, w, 4] arrays."""

    def __init__(self, hitables, materials, meshes=(), textures=(), camera=None, name=""):
        self.hitables = np.ascontiguousarray(hitables, dtype=hitable_dtype)
        self.materials = np.ascontiguousarray(materials, dtype=material_dtype)
        self.meshes = [np.ascontiguousarray(m, dtype=np.float32) for m in meshes]
        self.textures = [np.ascontiguousarray(t, dtype=np.uint8) for t in textures]
        self.camera = camera
        self.name = name

    def desc(self):
        """Build the rt_scene_desc; returns (desc, keepalive)."""
        n_m, n_t = len(self.meshes), len(self.textures)
        meshes = (RtMesh * max(n_m, 1))()
        for i, m in enumerate(self.meshes):
            assert m.ndim == 2 and m.shape[1] in (9, 15)
            meshes[i] = RtMesh(m.ctypes.data, m.shape[0], 1 if m.shape[1] == 15 else 0)
        texs = (RtTexture * max(n_t, 1))()
        for i, t in enumerate(self.textures):
            assert t.ndim == 3 and t.shape[2] == 4
            texs[i] = RtTexture(t.ctypes.data, t.shape[0], t.shape[1], t.shape[1] * 4)
        d = RtSceneDesc(
            self.hitables.ctypes.data, len(self.hitables), self.materials.ctypes.data, len(self.materials),
            meshes, n_m, texs, n_t,
        )
        return d, (meshes, texs, self)

    # ---- (de)serialisation for tests/golden -------------------------------------------------
    def save(self, path):
        arrs = {"hitables": self.hitables, "materials": self.materials}
        if self.camera is not None:
            arrs["camera"] = self.camera
        for i, m in enumerate(self.meshes):
            arrs[f"mesh{i}"] = m
        for i, t in enumerate(self.textures):
            arrs[f"tex{i}"] = t
        np.savez_compressed(path, **arrs)

    @staticmethod
    def load(path, name=""):
        z = np.load(path)
        meshes, texs = [], []
        i = 0
        while f"mesh{i}" in z:
            meshes.append(z[f"mesh{i}"])
            i += 1
        i = 0
        while f"tex{i}" in z:
            texs.append(z[f"tex{i}"])
            i += 1
        cam = z["camera"] if "camera" in z else None
        return SceneData(z["hitables"], z["materials"], meshes, texs, cam, name)
