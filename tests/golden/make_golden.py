"""Generates tests/golden/*.npz from the reference's OWN sources compiled for the host (oracle/_ref/libref_host.so,
built from /root/reference by oracle/Makefile).  Run in the authoring container only:  python tests/golden/make_golden.py
Each file holds, for one reference scene: the flattened scene description the reference's InitWorld produced, seeded
rays, the reference's HitableList::Hit results on them, and a small PathTracing image with the reference's XORWOW
streams.  tests/test_golden.py checks the oracle restatement (CPU) and the CUDA core (GPU) against them."""
import os
import sys

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, os.path.dirname(os.path.dirname(HERE)))
import checkers as ck  # noqa: E402

scenes = ck.pkg.scenes
BOX = {"cornell_box": (0, 555), "birthday": (0, 555), "spheres": (-6, 6), "bunny": (-0.1, 0.2)}
IMG = {"cornell_box": (18, 32, 3), "birthday": (18, 32, 2), "spheres": (18, 32, 2), "bunny": (9, 16, 1)}


def main():
    assert ck.have_ref(), "needs /root/reference"
    for name in ("cornell_box", "spheres", "birthday", "bunny"):
        faces = tex = None
        if name == "bunny":
            faces = scenes.bunny_mesh(3000)
        if name == "birthday":
            tex = scenes.earth_texture(32, 64)
        ref = ck.RefScene.named(name, 10086, faces, tex)
        d = ref.describe()
        cam = ref.camera()
        o = ck.OracleScene(d, cam)  # only for its primary-ray generator (pinned to _ref by test_oracle_vs_ref)
        lo, hi = BOX[name]
        rays = np.concatenate([o.primary_rays(27, 48), ck.random_rays(3000, 11, lo, hi)]).astype(np.float32)
        hits = ref.trace(rays)
        hits_finite = ref.trace(rays[:1500], 0.05 * (hi - lo), 0.6 * (hi - lo))
        h, w, spp = IMG[name]
        image = ref.render(h, w, spp, post=0)
        out = dict(hitables=d.hitables, materials=d.materials, camera=cam, rays=rays, hits=hits, hits_finite=hits_finite,
                   finite_range=np.array([0.05 * (hi - lo), 0.6 * (hi - lo)]), image=image, image_spp=np.int32(spp),
                   seed=np.uint64(10086),
                   state0_draws=np.int32(scenes.spheres(10086, 'rtl')[2] if name == 'spheres' else 0))
        if faces is not None:
            out["faces_caller_order"] = faces
            out["faces_sorted"] = d.meshes[0]
        if tex is not None:
            out["texture"] = tex
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), **out)
        print(name, {k: getattr(v, "shape", None) for k, v in out.items()})


if __name__ == "__main__":
    main()
