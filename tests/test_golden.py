"""Golden vectors generated from the reference's own sources (tests/golden/make_golden.py: oracle/_ref, i.e.
/root/reference compiled for the host) — they travel with the repo, so these tests run where the reference is absent.
CPU: the oracle restatement reproduces them bit-exactly.  GPU: the CUDA core (through the C-ABI) reproduces the hits."""
import os

import numpy as np
import pytest

import checkers as ck

abi = ck.pkg.abi
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
NAMES = ["cornell_box", "spheres", "birthday", "bunny"]


def _load(name):
    g = np.load(os.path.join(GOLD, name + ".npz"))
    meshes = [g["faces_caller_order"]] if "faces_caller_order" in g else []
    textures = [g["texture"]] if "texture" in g else []
    scene = abi.SceneData(g["hitables"], g["materials"], meshes, textures, camera=g["camera"], name=name)
    return g, scene


@pytest.mark.parametrize("name", NAMES)
def test_oracle_reproduces_golden(name):
    g, scene = _load(name)
    ck.set_pow_mode(1)  # the host build of the reference evaluates pow(x,2)/pow(x,.5) with glibc
    try:
        o = ck.OracleScene(scene, g["camera"])
        assert len(ck.hits_equal(o.trace(g["rays"], quirks=1), g["hits"])) == 0
        t0, t1 = g["finite_range"]
        assert len(ck.hits_equal(o.trace(g["rays"][:1500], float(t0), float(t1), quirks=1), g["hits_finite"])) == 0
        if "faces_sorted" in g:
            assert o.sorted_faces().tobytes() == g["faces_sorted"].tobytes()
        h, w, _ = g["image"].shape
        img, _ = o.render_reference(h, w, int(g["image_spp"]), int(g["seed"]), state0_draws=int(g["state0_draws"]),
                                    quirks=1, post=0)
        assert (img.view(np.uint32) == g["image"].view(np.uint32)).all()
    finally:
        ck.set_pow_mode(0)


@pytest.mark.gpu
@pytest.mark.parametrize("name", NAMES)
def test_cuda_core_reproduces_golden_hits(name):
    g, scene = _load(name)
    ctx = ck.pkg.Context(0)
    ctx.commit(scene)
    got = ctx.trace_closest(g["rays"])
    want = g["hits"]
    # ids / materials: always identical.  (golden = the reference's BVH incl. AABB::Hit quirks; the rays used here
    # are primary rays + random rays on which quirks and exact visibility agree for these scenes — asserted)
    ids = ("obj", "sub", "material")
    bad_ids = ck.hits_equal(got, want, ids)
    quirk = ck.hits_equal(ck.OracleScene(scene, g["camera"]).trace(g["rays"], quirks=0), want, ids)
    assert set(bad_ids.tolist()) <= set(quirk.tolist()), (len(bad_ids), len(quirk))
    ok = np.setdiff1d(np.arange(len(want)), quirk)
    has_spheres = (scene.hitables["type"] == abi.RT_HIT_SPHERE).any()
    if not has_spheres:
        assert len(ck.hits_equal(got[ok], want[ok], ("t", "u", "v", "normal"))) == 0
    else:
        # Sphere::Hit's pow(x,2)/pow(x,.5): glibc pow in the golden build vs x*x / sqrt on the device — last bits of t
        sph = scene.hitables["type"][np.maximum(want["obj"], 0)] == abi.RT_HIT_SPHERE
        nonsph = ok[~sph[ok]]
        assert len(ck.hits_equal(got[nonsph], want[nonsph], ("t", "normal"))) == 0
        s_ok = ok[sph[ok] & (want["obj"][ok] >= 0)]
        rel = np.abs(got["t"][s_ok] - want["t"][s_ok]) / np.abs(want["t"][s_ok])
        assert rel.max() <= 1e-12 and (rel > 0).mean() < 0.01, (rel.max(), (rel > 0).mean())
    t0, t1 = g["finite_range"]
    gotf = ctx.trace_closest(g["rays"][:1500], float(t0), float(t1))
    assert (gotf["obj"] == g["hits_finite"]["obj"]).mean() > 0.999
    ctx.close()
