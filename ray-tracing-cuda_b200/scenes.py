"""Host-side scene builders: the reference's four scenes (scenes/*.cu) and the synthetic triangle
soup of BASELINE.json, restated as plain arrays (SceneData) for the C-ABI.

Every builder cites the scene file it mirrors.  Geometry that the reference derives from libm
(rotateY -> cosf/sinf) or from curand draws (spheres layout) is recomputed here from the same
formulas; tests/test_scenes.py checks the result against the reference's own InitWorld run on
the host (tests/golden/scene_*.npz).

The model/texture files the reference loads (resources/bunny.obj, resources/earthmap.jpg) are not
in its tree (.gitignore:1-2); `bunny_mesh` and `earth_texture` synthesise stand-ins of the same
size.
"""
import math

import numpy as np

from .abi import (
    RT_HIT_MESH,
    RT_HIT_PARALLELOGRAM,
    RT_HIT_SKY,
    RT_HIT_SPHERE,
    RT_MAT_DIELECTRIC,
    RT_MAT_DIFFUSE_LIGHT,
    RT_MAT_LAMBERTIAN,
    RT_MAT_METAL,
    RT_MAT_SKY,
    SceneData,
    hitable_dtype,
    material_dtype,
)

f32 = np.float32


# --------------------------------------------------------------------------------------------
# curand XORWOW, subsequence 0, offset 0 (what states[0] is after CudaRandomInit, utils.cu:43-47)
# --------------------------------------------------------------------------------------------
class Xorwow:
    """curand_init(seed, 0, 0) + curand_uniform, restated from the published XORWOW generator
    (Marsaglia 2003) and curand's documented seeding; checked against curand_kernel.h in tests."""

    M = 0xFFFFFFFF

    def __init__(self, seed):
        s0 = (seed & self.M) ^ 0xAAD26B49
        s1 = ((seed >> 32) & self.M) ^ 0xF7DCEFDD
        t0 = (1099087573 * s0) & self.M
        t1 = (2591861531 * s1) & self.M
        self.d = (6615241 + t1 + t0) & self.M
        self.v = [
            (123456789 + t0) & self.M,
            362436069 ^ t0,
            (521288629 + t1) & self.M,
            88675123 ^ t1,
            (5783321 + t0) & self.M,
        ]
        self.draws = 0

    def next_u32(self):
        v, M = self.v, self.M
        t = v[0] ^ (v[0] >> 2)
        v[0], v[1], v[2], v[3] = v[1], v[2], v[3], v[4]
        v[4] = (v[4] ^ ((v[4] << 4) & M)) ^ (t ^ ((t << 1) & M))
        self.d = (self.d + 362437) & M
        self.draws += 1
        return (v[4] + self.d) & M

    def uniform(self):
        """curand_uniform: (0, 1]"""
        x = self.next_u32()
        return f32(f32(x) * f32(2.3283064365386963e-10) + f32(1.1641532182693481e-10))

    def random_float(self, mn, mx):
        """CudaRandomFloat (utils.cuh:22-27): t * (max - min) + min in float"""
        t = self.uniform()
        return f32(f32(t * f32(f32(mx) - f32(mn))) + f32(mn))


# --------------------------------------------------------------------------------------------
# small helpers
# --------------------------------------------------------------------------------------------
def _mat(type_, color=(0, 0, 0), fuzz=0.0, ior=0.0, texture=-1):
    m = np.zeros((), dtype=material_dtype)
    m["type"], m["texture"], m["color"], m["fuzz"], m["ior"] = type_, texture, np.asarray(color, f32), fuzz, ior
    return m


def _hit(type_, material, p=None, radius=0.0, mesh=-1, group=-1):
    h = np.zeros((), dtype=hitable_dtype)
    h["type"], h["material"], h["radius"], h["mesh"], h["group"] = type_, material, radius, mesh, group
    if p is not None:
        p = np.asarray(p, dtype=f32)
        h["p"][: p.shape[0]] = p
    return h


def parallelogram(p3, material, group=-1):
    """Parallelogram ctor (parallelogram.cu:10-15): p[3] = p[1] + p[2] - p[0] in float."""
    p = np.asarray(p3, dtype=f32).reshape(3, 3)
    p4 = np.empty((4, 3), f32)
    p4[:3] = p
    p4[3] = (p[1] + p[2]).astype(f32) - p[0]
    return _hit(RT_HIT_PARALLELOGRAM, material, p4, group=group)


def _rotate_y(v, angle):
    """glm::rotateY (gtx/rotate_vector.inl:93-106), float arithmetic; cos/sin rounded to float."""
    c, s = f32(math.cos(float(angle))), f32(math.sin(float(angle)))
    v = np.asarray(v, f32)
    r = v.copy()
    r[0] = f32(v[0] * c) + f32(v[2] * s)
    r[2] = f32(f32(-v[0]) * s) + f32(v[2] * c)
    return r


def parallelepiped(lengths, material, transform, group):
    """Parallelepiped(lengths, material, transform) (parallelepiped.cu:34-55) as its six
    parallelograms in AddCorner order (parallelepiped.cu:25-32)."""
    lengths = np.asarray(lengths, f32)
    p = np.zeros((4, 3), f32)
    q = np.tile(lengths, (4, 1)).astype(f32)
    for i in range(1, 4):
        p[i, i - 1] = lengths[i - 1]
        q[i, i - 1] = 0
    p = np.stack([transform(x) for x in p]).astype(f32)
    q = np.stack([transform(x) for x in q]).astype(f32)
    out = []
    for corner in (p, q):
        for i in range(1, 4):
            x, y = i, (1 if i + 1 == 4 else i + 1)
            out.append(parallelogram([corner[0], corner[x], corner[y]], material, group))
    return out


class CameraSpec:
    """Arguments of the look-at Camera constructors (camera.cu:6-38)."""

    def __init__(self, position, look_at, up, fov, aspect, aperture=-1.0, focus_distance=1.0):
        self.position, self.look_at, self.up = (np.asarray(a, f32) for a in (position, look_at, up))
        self.fov, self.aspect, self.aperture, self.focus_distance = fov, aspect, aperture, focus_distance


ASPECT_16_9 = 1280.0 / 720.0  # every scene bakes double(WIDTH)/HEIGHT with WIDTH=1280, HEIGHT=720


# --------------------------------------------------------------------------------------------
# scenes/cornell_box.cu:35-69
# --------------------------------------------------------------------------------------------
def _cornell_walls(back_wall_first_point):
    mats = [
        _mat(RT_MAT_SKY),
        _mat(RT_MAT_LAMBERTIAN, (0.65, 0.05, 0.05)),
        _mat(RT_MAT_LAMBERTIAN, (0.12, 0.45, 0.15)),
        _mat(RT_MAT_DIFFUSE_LIGHT, (1, 1, 1)),
        _mat(RT_MAT_LAMBERTIAN, (0.73, 0.73, 0.73)),
    ]
    SKY, RED, GREEN, LIGHT, WHITE = range(5)
    pts = np.array(
        [
            [0, 0, 0], [0, 555, 0], [0, 0, 555],
            [555, 0, 0], [555, 555, 0], [555, 0, 555],
            [213, 554, 332], [213, 554, 227], [343, 554, 332],
            [0, 0, 0], [555, 0, 0], [0, 0, 555],
            back_wall_first_point, [0, 555, 555], [555, 0, 555],
            [0, 555, 0], [555, 555, 0], [0, 555, 555],
        ],
        dtype=f32,
    )
    hs = [_hit(RT_HIT_SKY, SKY)]
    for k, m in enumerate([RED, GREEN, LIGHT, WHITE, WHITE, WHITE]):
        hs.append(parallelogram(pts[3 * k : 3 * k + 3], m))
    return mats, hs, WHITE


def cornell_box():
    mats, hs, WHITE = _cornell_walls([0, 0, 555])
    pi = f32(math.pi)
    a1 = f32(f32(-pi) * f32(0.1))
    a2 = f32(pi / f32(12.0))
    hs += parallelepiped((165, 165, 165), WHITE, lambda p: _rotate_y(p, a1) + np.array([130, 0, 165], f32), 0)
    hs += parallelepiped((165, 330, 165), WHITE, lambda p: _rotate_y(p, a2) + np.array([265, 0, 295], f32), 1)
    cam = CameraSpec((278, 278, -800), (278, 278, 0), (0, 1, 0), math.pi * 2 / 9, ASPECT_16_9)
    return SceneData(np.stack(hs), np.stack(mats), name="cornell_box"), cam


# --------------------------------------------------------------------------------------------
# scenes/birthday.cu:42-74 (+ :76-89 texture upload)
# --------------------------------------------------------------------------------------------
def earth_texture(height=1024, width=2048, seed=2):
    """Stand-in for resources/earthmap.jpg: deterministic low-frequency RGBA pattern."""
    rng = np.random.RandomState(seed)
    v, u = np.meshgrid(np.linspace(0, 1, height, endpoint=False), np.linspace(0, 1, width, endpoint=False), indexing="ij")
    img = np.zeros((height, width, 4), np.float64)
    for c in range(3):
        acc = np.zeros_like(u)
        for k in range(1, 6):
            ph = rng.uniform(0, 2 * math.pi, 2)
            acc += np.sin(2 * math.pi * k * u + ph[0]) * np.cos(math.pi * k * v + ph[1]) / k
        img[..., c] = 0.5 + 0.22 * acc
    land = (np.sin(6 * math.pi * u) * np.sin(4 * math.pi * v) + 0.3 * np.sin(22 * math.pi * u + 1.0) * np.sin(14 * math.pi * v)) > 0.1
    img[..., 0] = np.where(land, img[..., 0] * 0.6 + 0.1, img[..., 0] * 0.2)
    img[..., 1] = np.where(land, img[..., 1] * 0.7 + 0.2, img[..., 1] * 0.4 + 0.1)
    img[..., 2] = np.where(land, img[..., 2] * 0.3, img[..., 2] * 0.5 + 0.45)
    img[..., 3] = 1.0
    return (np.clip(img, 0, 1) * 255).astype(np.uint8)


def birthday(texture=None):
    mats, hs, _ = _cornell_walls([555, 555, 555])
    mats.append(_mat(RT_MAT_LAMBERTIAN, texture=0))
    hs.append(_hit(RT_HIT_SPHERE, len(mats) - 1, [[278, 278, 0]], radius=100.0))
    tex = earth_texture() if texture is None else texture
    cam = CameraSpec((278, 278, -800), (278, 278, 0), (0, 1, 0), math.pi * 2 / 9, ASPECT_16_9)
    return SceneData(np.stack(hs), np.stack(mats), textures=[tex], name="birthday"), cam


# --------------------------------------------------------------------------------------------
# scenes/spheres.cu:34-81
# --------------------------------------------------------------------------------------------
def spheres(seed=10086, arg_order="rtl"):
    """Layout drawn from XORWOW(seed, subsequence 0) exactly as InitWorld does (spheres.cu:54-80).

    `arg_order`: evaluation order of the draws inside `vec3 center(a + rand, 0.2, b + rand)` and
    `vec3(rand, rand, rand)` — unspecified by C++; "rtl" is what g++ emits (host build of the
    reference), "ltr" what nvcc's device front end emits.  Returns (scene, camera, n_draws).
    """
    rng = Xorwow(seed)
    mats = [
        _mat(RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5)),
        _mat(RT_MAT_DIELECTRIC, (1, 1, 1), ior=1.5),
        _mat(RT_MAT_LAMBERTIAN, (0.4, 0.2, 0.1)),
        _mat(RT_MAT_METAL, (0.7, 0.6, 0.5), fuzz=0.0),
        _mat(RT_MAT_SKY),
    ]
    hs = [
        _hit(RT_HIT_SPHERE, 0, [[0, -1000, 0]], radius=1000.0),
        _hit(RT_HIT_SPHERE, 1, [[0, 1, 0]], radius=1.0),
        _hit(RT_HIT_SPHERE, 2, [[-4, 1, 0]], radius=1.0),
        _hit(RT_HIT_SPHERE, 3, [[4, 1, 0]], radius=1.0),
        _hit(RT_HIT_SKY, 4),
    ]
    for a in range(-11, 11):
        for b in range(-11, 11):
            choose = rng.random_float(0, 1)
            d0 = rng.random_float(0, 0.9)
            d1 = rng.random_float(0, 0.9)
            if arg_order == "rtl":
                d0, d1 = d1, d0
            center = np.array([f32(f32(a) + d0), f32(0.2), f32(f32(b) + d1)], f32)
            # `(center - vec3(4, 0.2, 0)).length() > 0.9` is glm's component count (3): always true
            c3 = [rng.random_float(0, 1) for _ in range(3)]
            if arg_order == "rtl":
                c3 = c3[::-1]
            if float(choose) < 0.8:
                mats.append(_mat(RT_MAT_LAMBERTIAN, c3))
            elif float(choose) < 0.95:
                fuzz = rng.random_float(0, 0.5)
                mats.append(_mat(RT_MAT_METAL, c3, fuzz=min(float(fuzz), 1.0)))
            else:
                mats.append(_mat(RT_MAT_DIELECTRIC, (1, 1, 1), ior=1.5))
            hs.append(_hit(RT_HIT_SPHERE, len(mats) - 1, [center], radius=0.2))
    cam = CameraSpec((13, 2, 3), (0, 0, 0), (0, 1, 0), math.pi * 1 / 9, ASPECT_16_9)
    return SceneData(np.stack(hs), np.stack(mats), name="spheres"), cam, rng.draws


# --------------------------------------------------------------------------------------------
# scenes/bunny.cu:44-62
# --------------------------------------------------------------------------------------------
def _value_noise3(p, seed, octaves=3):
    rng = np.random.RandomState(seed)
    out = np.zeros(len(p))
    amp, freq = 1.0, 2.0
    for _ in range(octaves):
        k = rng.uniform(-1, 1, (4, 3)) * freq
        ph = rng.uniform(0, 2 * math.pi, 4)
        out += amp * np.mean(np.sin(p @ k.T + ph), axis=1)
        amp *= 0.5
        freq *= 2.0
    return out


def bunny_mesh(n_faces=69451, seed=1):
    """Stand-in for resources/bunny.obj: an open displaced UV-sphere with exactly `n_faces`
    triangles fitted to the Stanford bunny's extents so the hard-coded camera frames it."""
    seg = 4
    while 2 * seg * (seg - 1) < n_faces:
        seg += 1
    rings = seg
    th = np.linspace(0, math.pi, rings + 1)[:, None]
    ph = np.linspace(0, 2 * math.pi, seg, endpoint=False)[None, :]
    dirs = np.stack([np.sin(th) * np.cos(ph), np.cos(th) * np.ones_like(ph), np.sin(th) * np.sin(ph)], -1).reshape(-1, 3)
    r = 1.0 + 0.25 * _value_noise3(dirs, seed)
    v = dirs * r[:, None]
    lo, hi = np.array([-0.095, 0.033, -0.062]), np.array([0.061, 0.187, 0.059])
    v = (v - v.min(0)) / (v.max(0) - v.min(0)) * (hi - lo) + lo
    v = v.astype(f32)
    idx = lambda i, j: i * seg + (j % seg)
    tris = []
    for i in range(rings):
        for j in range(seg):
            a, b, c, d = idx(i, j), idx(i, j + 1), idx(i + 1, j), idx(i + 1, j + 1)
            if i > 0:
                tris.append((a, c, b))
            if i < rings - 1:
                tris.append((b, c, d))
    tris = np.array(tris[:n_faces], np.int64)
    assert len(tris) == n_faces, (len(tris), n_faces)
    return v[tris].reshape(n_faces, 9).astype(f32)


def write_obj(path, faces9):
    """Write a Face<false> array as a Wavefront OBJ (what Model<false> loads, bunny.cu:65)."""
    f = np.asarray(faces9, f32).reshape(-1, 3)
    with open(path, "w") as fh:
        for v in f:
            fh.write("v %.9g %.9g %.9g\n" % (v[0], v[1], v[2]))
        for i in range(len(f) // 3):
            fh.write("f %d %d %d\n" % (3 * i + 1, 3 * i + 2, 3 * i + 3))


def bunny(faces9=None):
    mats = [_mat(RT_MAT_LAMBERTIAN, (0.12, 0.45, 0.15)), _mat(RT_MAT_SKY), _mat(RT_MAT_LAMBERTIAN, (1, 1, 1))]
    quad = np.array([[-0.025 - 0.5, 0.1 - 0.5, 2], [-0.025 + 0.5, 0.1 - 0.5, 2], [-0.025 - 0.5, 0.1 + 0.5, 2]], np.float64)
    hs = [parallelogram(quad.astype(f32), 0), _hit(RT_HIT_SKY, 1), _hit(RT_HIT_MESH, 2, mesh=0)]
    faces = bunny_mesh() if faces9 is None else faces9
    cam = CameraSpec((-0.025, 0.1, -0.5), (-0.025, 0.1, 0), (0, 1, 0), math.pi * 2 / 9, ASPECT_16_9)
    return SceneData(np.stack(hs), np.stack(mats), meshes=[faces], name="bunny"), cam


# --------------------------------------------------------------------------------------------
# BASELINE.json config 5: procedural triangle soup
# --------------------------------------------------------------------------------------------
def soup_mesh(n_tris, seed=3):
    rng = np.random.Generator(np.random.Philox(seed))
    s = 0.5 * n_tris ** (-1.0 / 3.0)
    centres = rng.uniform(-1, 1, (n_tris, 1, 3)).astype(f32)
    offs = rng.uniform(-s, s, (n_tris, 3, 3)).astype(f32)
    return (centres + offs).reshape(n_tris, 9).astype(f32)


def soup(n_tris, seed=3, aspect=3840.0 / 2160.0):
    mats = [_mat(RT_MAT_LAMBERTIAN, (0.7, 0.7, 0.7)), _mat(RT_MAT_SKY)]
    hs = [_hit(RT_HIT_MESH, 0, mesh=0), _hit(RT_HIT_SKY, 1)]
    cam = CameraSpec((0, 0, -4), (0, 0, 0), (0, 1, 0), math.radians(40.0), aspect)
    return SceneData(np.stack(hs), np.stack(mats), meshes=[soup_mesh(n_tris, seed)], name=f"soup{n_tris}"), cam


def by_name(name):
    """(SceneData, CameraSpec) for a BASELINE.json scene name."""
    if name == "cornell_box":
        return cornell_box()
    if name == "birthday":
        return birthday()
    if name == "spheres":
        s, c, _ = spheres()
        return s, c
    if name == "bunny":
        return bunny()
    if name.startswith("soup"):
        return soup(int(name[4:]))
    raise ValueError(name)
