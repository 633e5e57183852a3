"""Parity of the CUDA path (through the C-ABI of librtcore.so) against the CPU oracle.
Bit-exact for closest hits (t, ids, u, v, normal) and for rendered images (same keyed Philox
draws, same arithmetic); the one exception is sphere u,v (acosf/atan2f differ between CUDA's and
glibc's libm by design), so image-textured spheres are compared by mismatch fraction."""
import math

import os
import numpy as np
import pytest

import checkers as ck

pytestmark = pytest.mark.gpu

scenes = ck.pkg.scenes
abi = ck.pkg.abi
rt = ck.pkg


def _named(name):
    if name == "cornell_box":
        return scenes.cornell_box()
    if name == "spheres":
        s, c, _ = scenes.spheres(10086, "rtl")
        return s, c
    if name == "birthday":
        return scenes.birthday(scenes.earth_texture(64, 128))
    if name == "bunny9k":
        return scenes.bunny(scenes.bunny_mesh(9000))
    if name == "bunny":
        return scenes.bunny()
    if name == "soup100k":
        return scenes.soup(100000)
    raise ValueError(name)


_BOX = {"cornell_box": (0, 555), "birthday": (0, 555), "spheres": (-6, 6), "bunny9k": (-0.1, 0.2), "bunny": (-0.1, 0.2),
        "soup100k": (-1.2, 1.2)}


@pytest.fixture(scope="module")
def ctx():
    c = rt.Context(0)
    yield c
    c.close()


def _setup(ctx, name):
    s, c = _named(name)
    ctx.commit(s)
    cam = rt.core.camera_look_at(c)
    ctx.set_camera(cam)
    o = ck.OracleScene(s, cam)
    return s, cam, o


def _fields(name):
    # sphere u,v go through acosf/atan2f: excluded where spheres exist
    if name in ("spheres", "birthday"):
        return ("t", "obj", "sub", "normal", "material")
    return ("t", "obj", "sub", "u", "v", "normal", "material")


@pytest.mark.parametrize("name", ["cornell_box", "spheres", "birthday", "bunny9k", "bunny", "soup100k"])
def test_closest_hit_bit_exact(ctx, name):
    s, cam, o = _setup(ctx, name)
    H, W = (108, 192) if name in ("bunny", "soup100k") else (135, 240)
    rays = ctx.primary_rays(H, W)
    assert rays.tobytes() == o.primary_rays(H, W).tobytes()
    want = o.trace(rays, quirks=0)
    got = ctx.trace_closest(rays, mode=0)
    bad = ck.hits_equal(got, want, _fields(name))
    assert len(bad) == 0, (len(bad), got[bad[:3]], want[bad[:3]])
    # the reference's own BVH semantics (AABB::Hit quirks included) agree on primary rays from outside
    if name in ("bunny9k", "cornell_box"):
        assert len(ck.hits_equal(got, o.trace(rays, quirks=1), ("t", "obj", "sub"))) == 0
    lo, hi = _BOX[name]
    n = 30000 if name in ("bunny", "soup100k") else 100000
    rr = ck.random_rays(n, 7, lo, hi)
    want = o.trace(rr, quirks=0)
    got = ctx.trace_closest(rr, mode=0)
    bad = ck.hits_equal(got, want, _fields(name))
    assert len(bad) == 0, (len(bad), got[bad[:3]], want[bad[:3]])
    # finite range
    t0, t1 = (0.05, 0.12) if name.startswith("bunny") else ((0.5, 300.0) if lo == 0 else (0.5, 3.0))
    want = o.trace(rr[:20000], t0, t1, quirks=0)
    got = ctx.trace_closest(rr[:20000], t0, t1, mode=0)
    assert len(ck.hits_equal(got, want, _fields(name))) == 0


@pytest.mark.parametrize("name", ["cornell_box", "spheres", "bunny9k"])
def test_sequential_scan_mode_bit_exact(ctx, name):
    s, cam, o = _setup(ctx, name)
    lo, hi = _BOX[name]
    rr = ck.random_rays(20000, 9, lo, hi)
    want = o.trace(rr, quirks=0)
    got = ctx.trace_closest(rr, mode=1)
    assert len(ck.hits_equal(got, want, _fields(name))) == 0


@pytest.mark.parametrize("name", ["cornell_box", "spheres", "bunny9k", "soup100k"])
def test_degenerate_directions(ctx, name):
    """Rays with direction components exactly +-0 / denormal / 1e-30 and origins exactly on box planes: the
    slab tests (FFMA form, clamped reciprocal) must never cull a record the reference accepts.  Every
    traversal (the scene's default, binary if-if / while-while) against the oracle's exhaustive scan."""
    s, cam, o = _setup(ctx, name)
    lo, hi = _BOX[name]
    rr = ck.degenerate_rays(40000, 33, lo, hi)
    want = o.trace(rr, quirks=0)
    for mode in (0, 2, 3):
        got = ctx.trace_closest(rr, mode=mode)
        bad = ck.hits_equal(got, want, _fields(name))
        assert len(bad) == 0, (mode, len(bad), rr[bad[:3]], got[bad[:3]], want[bad[:3]])


@pytest.mark.parametrize("name", ["cornell_box", "spheres", "bunny9k", "soup100k"])
def test_traversal_variants_identical(ctx, name):
    """Every tree traversal of the render kernel (mode 2 binary if-if, 3 binary while-while; the box list has its own
    test) returns the oracle's hits."""
    s, cam, o = _setup(ctx, name)
    lo, hi = _BOX[name]
    rr = np.concatenate([ctx.primary_rays(54, 96), ck.random_rays(30000, 21, lo, hi)])
    want = o.trace(rr, quirks=0)
    for mode in (2, 3):
        got = ctx.trace_closest(rr, mode=mode)
        bad = ck.hits_equal(got, want, _fields(name))
        assert len(bad) == 0, (mode, len(bad), got[bad[:3]], want[bad[:3]])


@pytest.mark.parametrize("name", ["cornell_box", "birthday", "mixed30"])
def test_box_list_traversal(ctx, name):
    """Mode 4: the box-list traversal of scenes with at most 32 primitives (every box tested by every lane, nearest
    box first, the rest re-tested against the closest hit) returns the oracle's hits on random, primary and
    degenerate rays; larger scenes refuse the mode."""
    if name == "mixed30":
        rng = np.random.RandomState(8)
        mats = [scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5)), scenes._mat(abi.RT_MAT_METAL, (0.7, 0.6, 0.5), fuzz=0.3),
                scenes._mat(abi.RT_MAT_SKY)]
        hs = []
        for i in range(30):
            k = i % 3
            if k == 0:
                hs.append(scenes._hit(abi.RT_HIT_SPHERE, i % 2, [rng.uniform(-3, 3, 3)], radius=float(rng.uniform(0.1, 1.0))))
            elif k == 1:
                hs.append(scenes._hit(abi.RT_HIT_TRIANGLE, i % 2, rng.uniform(-3, 3, (3, 3))))
            else:
                hs.append(scenes.parallelogram(rng.uniform(-3, 3, (3, 3)), i % 2))
        hs.append(scenes._hit(abi.RT_HIT_SKY, 2))
        s = abi.SceneData(np.stack(hs), np.stack(mats))
        ctx.commit(s)
        o = ck.OracleScene(s)
        lo, hi, fields = -4, 4, ("t", "obj", "sub", "normal", "material")
        rr = ck.random_rays(40000, 5, lo, hi)
    else:
        s, cam, o = _setup(ctx, name)
        lo, hi = _BOX[name]
        fields = _fields(name)
        rr = np.concatenate([ctx.primary_rays(54, 96), ck.random_rays(40000, 5, lo, hi)])
    rr = np.concatenate([rr, ck.degenerate_rays(20000, 34, lo, hi)])
    want = o.trace(rr, quirks=0)
    got = ctx.trace_closest(rr, mode=4)
    bad = ck.hits_equal(got, want, fields)
    assert len(bad) == 0, (len(bad), rr[bad[:3]], got[bad[:3]], want[bad[:3]])
    # ranges other than the render default
    for rng_ in ((0.0, 1e9), (1e-3, 300.0)):
        got = ctx.trace_closest(rr[:20000], *rng_, mode=4)
        assert len(ck.hits_equal(got, o.trace(rr[:20000], *rng_, quirks=0), fields)) == 0, rng_
    # a scene with more than 32 primitives has no box list
    _setup(ctx, "spheres")
    with pytest.raises(Exception):
        ctx.trace_closest(rr[:10], mode=4)


def test_mixed_scene_all_primitive_types(ctx):
    rng = np.random.RandomState(3)
    mats = [scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5)), scenes._mat(abi.RT_MAT_METAL, (0.7, 0.6, 0.5), fuzz=0.3),
            scenes._mat(abi.RT_MAT_DIELECTRIC, (1, 1, 1), ior=1.5), scenes._mat(abi.RT_MAT_DIFFUSE_LIGHT, (2, 2, 2)),
            scenes._mat(abi.RT_MAT_SKY)]
    hs = []
    for i in range(60):
        k = i % 3
        if k == 0:
            hs.append(scenes._hit(abi.RT_HIT_SPHERE, i % 4, [rng.uniform(-3, 3, 3)], radius=float(rng.uniform(0.1, 1.0))))
        elif k == 1:
            hs.append(scenes._hit(abi.RT_HIT_TRIANGLE, i % 4, rng.uniform(-3, 3, (3, 3))))
        else:
            hs.append(scenes.parallelogram(rng.uniform(-3, 3, (3, 3)), i % 4))
        if i == 20:
            hs.append(scenes._hit(abi.RT_HIT_SKY, 4))
    mesh = scenes.soup_mesh(3000, seed=5) * np.float32(2.0)
    uvmesh = np.concatenate([scenes.soup_mesh(500, seed=6) * np.float32(2.5), rng.uniform(0, 1, (500, 6)).astype(np.float32)], 1)
    hs.append(scenes._hit(abi.RT_HIT_MESH, 0, mesh=0))
    hs.append(scenes._hit(abi.RT_HIT_MESH, 1, mesh=1))
    s = abi.SceneData(np.stack(hs), np.stack(mats), meshes=[mesh, uvmesh])
    ctx.commit(s)
    o = ck.OracleScene(s)
    rays = ck.random_rays(100000, 5, -4, 4)
    fields = ("t", "obj", "sub", "normal", "material")
    for rng_ in ((1e-3, float("inf")), (0.5, 3.0)):
        want = o.trace(rays, *rng_, quirks=0)
        got = ctx.trace_closest(rays, *rng_, mode=0)
        bad = ck.hits_equal(got, want, fields)
        assert len(bad) == 0, (len(bad), got[bad[:3]], want[bad[:3]])
        nonsphere = s.hitables["type"][np.maximum(want["obj"], 0)] != abi.RT_HIT_SPHERE
        assert len(ck.hits_equal(got[nonsphere], want[nonsphere], ("u", "v"))) == 0
    got1 = ctx.trace_closest(rays[:20000], mode=1)
    assert len(ck.hits_equal(got1, o.trace(rays[:20000], quirks=0), fields)) == 0


def test_degenerate_inputs(ctx):
    # empty scene, sky only, a single primitive, zero rays
    mats = np.stack([scenes._mat(abi.RT_MAT_SKY), scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5))])
    rays = ck.random_rays(1000, 1, -1, 1)
    empty = abi.SceneData(np.zeros(0, abi.hitable_dtype), mats)
    ctx.commit(empty)
    assert (ctx.trace_closest(rays)["obj"] == -1).all()
    assert len(ctx.trace_closest(rays[:0])) == 0
    sky = abi.SceneData(np.stack([scenes._hit(abi.RT_HIT_SKY, 0)]), mats)
    ctx.commit(sky)
    h = ctx.trace_closest(rays)
    assert (h["obj"] == 0).all() and (h["t"] == 1e9).all()
    assert (ctx.trace_closest(rays, 1e-3, 5.0)["obj"] == -1).all()
    one = abi.SceneData(np.stack([scenes._hit(abi.RT_HIT_TRIANGLE, 1, [[-5, -5, 2], [5, -5, 2], [0, 5, 2]])]), mats)
    ctx.commit(one)
    o = ck.OracleScene(one)
    assert len(ck.hits_equal(ctx.trace_closest(rays), o.trace(rays, quirks=0))) == 0


@pytest.mark.parametrize("name,hw,spp", [("cornell_box", (72, 128), 16), ("spheres", (54, 96), 8), ("bunny9k", (54, 96), 4)])
def test_render_bit_exact_vs_oracle(ctx, name, hw, spp):
    import torch
    s, cam, o = _setup(ctx, name)
    H, W = hw
    want, cnt = o.render_product(H, W, 0, spp, 1234)
    # wavefront: the single shade queue (default), material-sorted queues (SORT) and a reordered extend queue (RAYSORT)
    # give the same sums
    for flags in (abi.RT_RENDER_FUSED, abi.RT_RENDER_WAVEFRONT, abi.RT_RENDER_WAVEFRONT | abi.RT_RENDER_SORT,
                  abi.RT_RENDER_WAVEFRONT | abi.RT_RENDER_RAYSORT):
        got, _, st = ctx.render_host(H, W, spp, seed=1234, flags=flags)
        diff = np.nonzero((got.view(np.uint32) != want.view(np.uint32)).any(axis=2))
        assert len(diff[0]) == 0, (flags, len(diff[0]), got[diff][:3], want[diff][:3])
        assert (st.samples, st.rays, st.rays_ref_equiv) == tuple(int(x) for x in cnt)
    # the float interface adds the (rounded) sum of each call: the oracle run the same way
    a, _ = o.render_product(H, W, 0, spp // 2, 1234)
    a, _ = o.render_product(H, W, spp // 2, spp - spp // 2, 1234, accum=a)
    d = torch.zeros(H, W, 3, device="cuda")
    ctx.render_device(d.data_ptr(), H, W, spp // 2, 0, seed=1234)
    ctx.render_device(d.data_ptr(), H, W, spp - spp // 2, spp // 2, seed=1234)
    torch.cuda.synchronize()
    assert (d.cpu().numpy().view(np.uint32) == a.view(np.uint32)).all()
    # the exact sums: any split of the sample range, either strategy, equals the oracle's one-shot sums
    fx, _ = o.render_product_fixed(H, W, 0, spp, 1234)
    q = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    ctx.render_fixed_device(q.data_ptr(), H, W, spp // 3, 0, seed=1234, flags=abi.RT_RENDER_CLEAR)
    ctx.render_fixed_device(q.data_ptr(), H, W, spp - spp // 3, spp // 3, seed=1234, flags=abi.RT_RENDER_WAVEFRONT)
    torch.cuda.synchronize()
    assert np.array_equal(q.cpu().numpy(), fx)
    acc = torch.zeros(H, W, 3, device="cuda")
    ctx.resolve_fixed_device(q.data_ptr(), H, W, spp, d_accum_ptr=acc.data_ptr())
    torch.cuda.synchronize()
    assert np.array_equal(acc.cpu().numpy().view(np.uint32), want.view(np.uint32))


@pytest.mark.parametrize("name,hw,spp", [("cornell_box", (72, 128), 8), ("spheres", (54, 96), 8), ("bunny9k", (54, 96), 4)])
def test_render_kernel_variants_bit_exact(ctx, name, hw, spp, monkeypatch):
    """Every traversal the render kernel is compiled for (RT_TRAV, read at commit: 0 binary if-if at 64 registers,
    1 binary while-while, 20 box list where the scene has one) renders the oracle's image."""
    s, cam = _named(name)
    camera = rt.core.camera_look_at(cam)
    o = ck.OracleScene(s, camera)
    H, W = hw
    want, cnt = o.render_product(H, W, 0, spp, 99)
    for trav in (0, 1, 20):
        monkeypatch.setenv("RT_TRAV", str(trav))
        ctx.commit(s)
        ctx.set_camera(camera)
        got, _, st = ctx.render_host(H, W, spp, seed=99, flags=abi.RT_RENDER_FUSED)
        assert (got.view(np.uint32) == want.view(np.uint32)).all(), trav
        assert (st.samples, st.rays, st.rays_ref_equiv) == tuple(int(x) for x in cnt), trav
    monkeypatch.delenv("RT_TRAV")
    ctx.commit(s)


@pytest.mark.parametrize("name,hw,spp", [("spheres", (54, 96), 8), ("cornell_box", (67, 125), 7), ("bunny9k", (54, 96), 37)])
def test_work_distribution_does_not_change_the_image(ctx, name, hw, spp, monkeypatch):
    """The path kernel hands out (pixel, sample chunk) sub-units from (tile, chunk group) items drawn from a global
    counter; which warp renders what is a race by design.  The exact sums do not depend on it: forced chunk sizes of
    1, 3 (ragged against 7, 8 and 37 samples) and 16, items of 1 and 5 chunks, an image whose last tile row / column
    is ragged — always the oracle's sums and counters."""
    import torch
    s, cam, o = _setup(ctx, name)
    H, W = hw
    want, cnt = o.render_product_fixed(H, W, 0, spp, 77)
    q = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    for chunk, item in ((None, None), ("1", "1"), ("3", "5"), ("16", "1"), ("2", "16")):
        if chunk:
            monkeypatch.setenv("RT_CHUNK", chunk)
            monkeypatch.setenv("RT_ITEM_CHUNKS", item)
        st = ctx.render_fixed_device(q.data_ptr(), H, W, spp, 0, seed=77, flags=abi.RT_RENDER_CLEAR)
        torch.cuda.synchronize()
        assert np.array_equal(q.cpu().numpy(), want), (chunk, item)
        assert (st.samples, st.rays, st.rays_ref_equiv) == tuple(int(x) for x in cnt)
        assert st.kernel_launches == 1 and st.work_items > 0
        if chunk:
            assert st.chunk_samples == min(int(chunk), spp)
    monkeypatch.delenv("RT_CHUNK", raising=False)
    monkeypatch.delenv("RT_ITEM_CHUNKS", raising=False)
    # explicit phase lists (chunk x chunks-per-item : samples): ragged last items inside a phase, four phases, a phase
    # list that stops short of / overshoots the sample range
    for plan in ("4x2:6,3x3:1,2x1", "16x4", "2x3:4,1x1:1,5x2:2,3x1", "1x64:3,4x1", "8x1:1000"):
        monkeypatch.setenv("RT_PLAN", plan)
        st = ctx.render_fixed_device(q.data_ptr(), H, W, spp, 0, seed=77, flags=abi.RT_RENDER_CLEAR)
        torch.cuda.synchronize()
        assert np.array_equal(q.cpu().numpy(), want), plan
        assert (st.samples, st.rays, st.rays_ref_equiv) == tuple(int(x) for x in cnt), plan
    monkeypatch.delenv("RT_PLAN")


def _mixed_small(n, seed):
    """n primitives of all three kinds + a sky in the middle of the list (records: n/3 spheres, n/3 triangles, 2n/3 quad halves)."""
    rng = np.random.RandomState(seed)
    mats = [scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5)), scenes._mat(abi.RT_MAT_METAL, (0.7, 0.6, 0.5), fuzz=0.3),
            scenes._mat(abi.RT_MAT_SKY)]
    hs = []
    for i in range(n):
        k = i % 3
        if k == 0:
            hs.append(scenes._hit(abi.RT_HIT_SPHERE, i % 2, [rng.uniform(-3, 3, 3)], radius=float(rng.uniform(0.05, 0.5))))
        elif k == 1:
            hs.append(scenes._hit(abi.RT_HIT_TRIANGLE, i % 2, rng.uniform(-3, 3, (3, 3))))
        else:
            hs.append(scenes.parallelogram(rng.uniform(-3, 3, (3, 3)), i % 2))
        if i == n // 2:
            hs.append(scenes._hit(abi.RT_HIT_SKY, 2))
    return abi.SceneData(np.stack(hs), np.stack(mats))


@pytest.mark.parametrize("name", ["cornell_box", "birthday", "spheres", "one", "two", "mixed33", "mixed1024", "mixed1025",
                                  "dup_keys"])
def test_small_scene_builder_identical(ctx, name, monkeypatch):
    """Scenes of at most 1024 primitives / 2048 records are built by ONE launch (k_build_small: every step of the
    LBVH build in one block); RT_BUILD_SMALL=0 sends them through the step-per-launch path of the large scenes.  Same
    per-element bodies, same arithmetic, a stable sort in both: the two trees must be the same tree — node count, depth
    and SAH cost equal bit for bit, the same hits, and the same number of box and record tests on the same rays (a
    different tree would visit different nodes).  1025 primitives take the step-per-launch path either way."""
    cam = None
    if name in ("cornell_box", "birthday", "spheres"):
        s, cam = _named(name)
    elif name == "one":
        s = abi.SceneData(np.stack([scenes._hit(abi.RT_HIT_TRIANGLE, 0, [[-5, -5, 2], [5, -5, 2], [0, 5, 2]])]),
                          np.stack([scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5))]))
    elif name == "two":
        s = _mixed_small(2, 4)
    elif name == "dup_keys":  # coincident primitives: equal Morton keys, the order is the sort's stability
        rng = np.random.RandomState(12)
        tri = rng.uniform(-2, 2, (3, 3))
        hs = [scenes._hit(abi.RT_HIT_TRIANGLE, 0, tri + (0.0 if i % 4 else 0.25 * (i // 4))) for i in range(40)]
        s = abi.SceneData(np.stack(hs), np.stack([scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5))]))
    else:
        s = _mixed_small(int(name[5:]), 17)
    lo, hi = _BOX.get(name, (-4, 4))
    rr = ck.random_rays(30000, 9, lo, hi)
    res = {}
    for small in ("0", "1"):
        monkeypatch.setenv("RT_BUILD_SMALL", small)
        bs = ctx.commit(s)
        if cam is not None:
            ctx.set_camera(rt.core.camera_look_at(cam))
        else:
            ctx.set_camera(rt.core.camera_look_at(scenes.cornell_box()[1]))
        hits = ctx.trace_closest(rr, mode=3)
        hits_default = ctx.trace_closest(rr, mode=0)
        _, _, st = ctx.render_host(27, 48, 2, seed=5, flags=abi.RT_RENDER_FUSED | abi.RT_RENDER_PROFILE)
        res[small] = ((bs.n_prims, bs.n_nodes, bs.max_depth, np.float32(bs.sah_cost).view(np.uint32).item(), bs.n_spheres,
                       bs.n_quads, bs.n_tris), hits, hits_default, (st.rays, st.node_visits, st.prim_tests))
    monkeypatch.delenv("RT_BUILD_SMALL")
    a, b = res["0"], res["1"]
    assert a[0] == b[0], (a[0], b[0])
    assert a[1].tobytes() == b[1].tobytes() and a[2].tobytes() == b[2].tobytes()
    assert a[3] == b[3], (a[3], b[3])
    o = ck.OracleScene(s)
    want = o.trace(rr, quirks=0)
    fields = ("t", "obj", "sub", "normal", "material")
    assert len(ck.hits_equal(b[1], want, fields)) == 0 and len(ck.hits_equal(b[2], want, fields)) == 0


def test_render_textured_sphere_close(ctx):
    s, cam, o = _setup(ctx, "birthday")
    H, W, spp = 54, 96, 8
    want, _ = o.render_product(H, W, 0, spp, 5)
    got, _, _ = ctx.render_host(H, W, spp, seed=5)
    frac = np.mean((got.view(np.uint32) != want.view(np.uint32)).any(axis=2))
    assert frac < 0.02, frac  # texel flips where CUDA acosf/atan2f differ from glibc by an ulp
    assert np.abs(got - want).sum() / np.abs(want).sum() < 5e-3


def test_max_depth_and_resolve(ctx):
    s, cam, o = _setup(ctx, "cornell_box")
    H, W, spp = 36, 64, 4
    for depth in (0, 1, 3, 50):
        want, cnt = o.render_product(H, W, 0, spp, 3, depth_limit=depth)
        got, rgb, st = ctx.render_host(H, W, spp, seed=3, max_depth=depth, want_rgb=True)
        assert (got.view(np.uint32) == want.view(np.uint32)).all(), depth
        assert st.rays_ref_equiv == int(cnt[2])
        ref_rgb = np.sqrt(np.clip(want / np.float32(spp), 0, 1)).astype(np.float32)
        assert np.abs(rgb - ref_rgb).max() <= 1e-6


def test_traversals_agree_at_full_size(ctx):
    """A 1 M-triangle soup (the oracle would need minutes): the binary while-while loop returns exactly the hits of the
    binary if-if loop (shared-memory stack, the product path at this size) on 400 k primary and random rays, and on rays
    scaled far away from the scene box."""
    s, cam = scenes.soup(1000000)
    ctx.commit(s)
    ctx.set_camera(rt.core.camera_look_at(cam))
    rr = np.concatenate([ctx.primary_rays(270, 480), ck.random_rays(270000, 33, -1.2, 1.2)])
    far = rr[:50000].copy()
    far[:, :3] *= np.float32(50.0)  # origins far outside the scene box: large |o| / |d| against small boxes
    rr = np.concatenate([rr, far])
    want = ctx.trace_closest(rr, mode=2)
    assert (want["obj"] >= 0).mean() > 0.3
    for mode in (3, 0):
        got = ctx.trace_closest(rr, mode=mode)
        bad = ck.hits_equal(got, want)
        assert len(bad) == 0, (mode, len(bad), rr[bad[:3]], got[bad[:3]], want[bad[:3]])


def test_full_size_properties(ctx):
    """BASELINE sizes through size-independent properties: additivity over sample ranges and
    strategy independence (path kernel == wavefront) of the exact sums at 1920x1080 on the full bunny."""
    import torch
    s, cam, o = _setup(ctx, "bunny")
    H, W = 1080, 1920
    a = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    b = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    ctx.render_fixed_device(a.data_ptr(), H, W, 2, 0, seed=11, flags=abi.RT_RENDER_FUSED)
    ctx.render_fixed_device(b.data_ptr(), H, W, 2, 0, seed=11, flags=abi.RT_RENDER_WAVEFRONT)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    c = torch.zeros(H, W, 3, device="cuda", dtype=torch.int64)
    ctx.render_fixed_device(c.data_ptr(), H, W, 1, 1, seed=11)
    ctx.render_fixed_device(c.data_ptr(), H, W, 1, 0, seed=11)
    torch.cuda.synchronize()
    assert torch.equal(a, c)


def test_rays_through_a_parallelogram_diagonal(ctx):
    """Parallelogram::Hit answers with its FIRST triangle that hits (parallelogram.cu:25-41); on the shared diagonal
    both hit, and when the second is closer the answer depends on what HitableList::Hit found before it
    (hitable_list.cu:11-21).  Rays aimed at the diagonal, with other objects listed before and after the parallelogram,
    in front of it, behind it and coplanar with it: every traversal returns the oracle's hits (resolve_quad_order; the
    sequential scan only for what it cannot settle), on a scene with a mesh too (where that scan would test every face)."""
    rng = np.random.RandomState(12)
    mats = np.stack([scenes._mat(abi.RT_MAT_LAMBERTIAN, (0.5, 0.5, 0.5)), scenes._mat(abi.RT_MAT_SKY)])
    quad = np.array([[0, 0, 0], [1, 0, 0], [0, 1, 0]], np.float32)

    def rays_at_diagonal(n):
        s = rng.uniform(0.02, 0.98, n).astype(np.float32)
        target = np.stack([s, np.float32(1) - s, np.zeros(n, np.float32)], 1)  # x + y = 1: the edge p1-p2
        origin = (target + rng.normal(size=(n, 3)) * 0.7 + np.array([0, 0, 2.0])).astype(np.float32)
        origin[n // 2:, 2] *= -1  # from both sides
        d = (target - origin).astype(np.float32)
        dot = (d[:, 0] * d[:, 0] + d[:, 1] * d[:, 1]).astype(np.float32) + (d[:, 2] * d[:, 2]).astype(np.float32)
        d = (d * (np.float32(1) / np.sqrt(dot))[:, None]).astype(np.float32)
        return np.concatenate([origin, d], 1).astype(np.float32)

    def shifted(dz):
        return quad + np.array([0, 0, dz], np.float32)

    mesh = (scenes.soup_mesh(5000, seed=9) * np.float32(0.8) + np.array([0.5, 0.5, -1.5], np.float32).repeat(3)).astype(np.float32)
    layouts = {
        "alone": [scenes.parallelogram(quad, 0), scenes._hit(abi.RT_HIT_SKY, 1)],
        "before_in_front": [scenes.parallelogram(shifted(0.5), 0), scenes.parallelogram(quad, 0), scenes._hit(abi.RT_HIT_SKY, 1)],
        "after_in_front": [scenes._hit(abi.RT_HIT_SKY, 1), scenes.parallelogram(quad, 0), scenes.parallelogram(shifted(0.5), 0)],
        "coplanar_before_after": [scenes.parallelogram(quad, 0), scenes.parallelogram(quad, 0), scenes.parallelogram(quad[[1, 0, 2]], 0)],
        "with_mesh": [scenes._hit(abi.RT_HIT_MESH, 0, mesh=0), scenes.parallelogram(quad, 0), scenes._hit(abi.RT_HIT_SKY, 1),
                      scenes.parallelogram(shifted(-0.25), 0)],
    }
    rays = rays_at_diagonal(60000)
    n_both = 0
    for name, hs in layouts.items():
        s = abi.SceneData(np.stack(hs), mats, meshes=[mesh] if name == "with_mesh" else [])
        ctx.commit(s)
        o = ck.OracleScene(s)
        want = o.trace(rays, quirks=0)
        for mode in (0, 1, 2, 3) + ((4,) if name != "with_mesh" else ()):
            got = ctx.trace_closest(rays, mode=mode)
            bad = ck.hits_equal(got, want, ("t", "obj", "sub", "normal", "material"))
            assert len(bad) == 0, (name, mode, len(bad), rays[bad[:2]], got[bad[:2]], want[bad[:2]])
        n_both += int((want["sub"] == 1).sum())
    assert n_both > 1000  # the second triangle does get reported
