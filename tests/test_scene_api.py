"""The reference's own scenes/*.cu, compiled UNCHANGED against ray-tracing-cuda_b200/scene_api (binaries built
here by `make -C ray-tracing-cuda_b200/scene_api`, they travel to the GPU box), run end to end:
InitWorld on the device -> flatten -> rt_scene_commit -> rt_render -> resolve -> image file, and are compared
with the same scene described through the Python bindings (bit-exact where the geometry is bit-reproducible)."""
import importlib
import json
import os
import subprocess
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
API = os.path.join(ROOT, "ray-tracing-cuda_b200", "scene_api")
BUILD = os.path.join(API, "build")
SCENES = ["spheres", "cornell_box", "bunny", "birthday"]


def _have_reference():
    return os.path.exists("/root/reference/scenes/spheres.cu")


@pytest.mark.skipif(not _have_reference(), reason="reference sources are only mounted in the authoring container")
def test_reference_scenes_compile_unchanged():
    """Acceptance test of the drop-in boundary: the four scene files build byte-for-byte unchanged."""
    subprocess.run(["make", "-s", "-j4", "-C", API], check=True)
    for s in SCENES:
        assert os.path.exists(os.path.join(BUILD, s)), s
    # nothing of the reference is compiled into them except scenes/<name>.cu itself
    out = subprocess.run(["make", "-n", "-B", "-C", API], capture_output=True, text=True).stdout
    assert "/root/reference/ray-tracing-cuda" not in out


def test_rtrun_launcher_sets_ranks(tmp_path):
    rtrun = os.path.join(BUILD, "rtrun")
    if not os.path.exists(rtrun):
        subprocess.run(["make", "-s", "-C", API, "build/rtrun"], check=True)
    out = subprocess.run([rtrun, "-n", "3", "sh", "-c", "echo $RT_RANK/$RT_WORLD_SIZE/$RT_JOB_ID"], capture_output=True,
                         text=True, check=True).stdout.split()
    assert sorted(x.split("/")[0] for x in out) == ["0", "1", "2"]
    assert {x.split("/")[1] for x in out} == {"3"} and len({x.split("/")[2] for x in out}) == 1
    assert subprocess.run([rtrun, "-n", "2", "sh", "-c", "exit $RT_RANK"]).returncode == 1


def _read_ppm(path):
    with open(path, "rb") as f:
        assert f.readline().strip() == b"P6"
        w, h = map(int, f.readline().split())
        assert f.readline().strip() == b"255"
        return np.frombuffer(f.read(), np.uint8).reshape(h, w, 3)


def _run_scene(name, tmp_path, H, W, spp, ranks=1):
    rt = importlib.import_module("ray-tracing-cuda_b200")
    res = tmp_path / "resources"
    if not res.exists():
        sys.path.insert(0, os.path.join(ROOT, "tools"))
        import make_resources
        make_resources.main(str(res), n_faces=9000, tex_hw=(64, 128))
    out, stats = tmp_path / f"{name}.ppm", tmp_path / f"{name}.json"
    env = dict(os.environ, RT_HEIGHT=str(H), RT_WIDTH=str(W), RT_SPP=str(spp), RT_OUT=str(out), RT_STATS=str(stats),
               RT_SEED="5", RT_LOG="0")
    cmd = [os.path.join(BUILD, name)]
    if ranks > 1:
        cmd = [os.path.join(BUILD, "rtrun"), "-n", str(ranks)] + cmd
    p = subprocess.run(cmd, cwd=tmp_path, env=env, capture_output=True, text=True, timeout=300)
    assert p.returncode == 0, p.stderr[-2000:]
    st = json.load(open(str(stats) + (".0" if ranks > 1 else "")))
    return rt, _read_ppm(out), st


def _python_render(rt, scene, cam, H, W, spp):
    ctx = rt.Context(0)
    ctx.commit(scene)
    ctx.set_camera(rt.core.camera_look_at(cam))
    acc, rgb, st = ctx.render_host(H, W, spp, seed=5, want_rgb=True)
    ctx.close()
    return (rgb * np.float32(255)).astype(np.uint8), st


@pytest.mark.gpu
@pytest.mark.parametrize("name", SCENES)
def test_scene_binary_matches_python_api(name, tmp_path):
    if not os.path.exists(os.path.join(BUILD, name)):
        pytest.skip("scene binaries not built (needs the reference's scenes/*.cu at build time)")
    H, W, spp = 90, 160, 8
    rt, img, st = _run_scene(name, tmp_path, H, W, spp)
    assert img.shape == (H, W, 3) and st["height"] == H and st["spp"] == spp and st["rays"] > H * W * spp
    sc = rt.scenes
    if name == "spheres":
        # layout drawn from states[0] inside the user kernel; argument evaluation order is the compiler's
        cands = [sc.spheres(10086, o)[:2] for o in ("ltr", "rtl")]
    elif name == "cornell_box":
        cands = [sc.cornell_box()]
    elif name == "bunny":
        cands = [sc.bunny(sc.bunny_mesh(9000))]
    else:
        from PIL import Image
        tex = np.asarray(Image.open(tmp_path / "resources" / "earthmap.jpg").convert("RGBA"))
        cands = [sc.birthday(tex)]
    best = None
    for scene, cam in cands:
        want, pst = _python_render(rt, scene, cam, H, W, spp)
        diff = np.abs(want.astype(int) - img.astype(int))
        frac = float((diff.max(axis=2) > 0).mean())
        if best is None or frac < best[0]:
            best = (frac, float(diff.mean()), pst)
    frac, mean_diff, pst = best
    if name in ("spheres", "bunny"):
        # geometry is bit-reproducible (float adds of curand draws / raw OBJ floats): identical images
        assert frac == 0.0, (frac, mean_diff)
        assert st["rays"] == pst.rays
    else:
        # rotateY goes through device cosf/sinf (cornell) and the JPEG decoder differs (birthday): statistical
        assert mean_diff < 1.5 and frac < 0.6, (frac, mean_diff)


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["spheres", "cornell_box"])
def test_two_ranks_render_the_one_rank_image(name, tmp_path):
    """DistributedMain (spheres) AND Main (cornell_box: the reference's Main is single-GPU, utils.cu:132-179) under
    `rtrun -n 2`: each rank renders its GetWorkload share of the GLOBAL sample indices into exact 64-bit sums, one
    ncclReduce (rt_reduce_fixed) adds them at rank 0, rank 0 writes the image.  Integer sums do not depend on the
    partition: the 2-rank image IS the 1-rank image, byte for byte."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    if not os.path.exists(os.path.join(BUILD, name)):
        pytest.skip("scene binaries not built")
    H, W, spp = 90, 160, 9  # odd: ranks get 5 and 4 samples (GetWorkload, utils.cu:111-113)
    _, one, st1 = _run_scene(name, tmp_path, H, W, spp, ranks=1)
    _, two, st2 = _run_scene(name, tmp_path, H, W, spp, ranks=2)
    assert st2["world_size"] == 2 and st2["local_spp"] == 5
    assert st2["rays"] < st1["rays"]  # rank 0 traced only its share
    assert np.array_equal(one, two)


_OBJ = """# quads, a pentagon, every corner syntax, negative indices, two materials (one textured, revisited)
mtllib a.mtl
v 0 0 0
v 1 0 0
v 1 1 0
v 0 1 0
v 0.5 1.5 0.25
vt 0 0
vt 1 0
vt 1 1
vt 0 1
vn 0 0 1
usemtl red
f 1/1 2/2 3/3 4/4
usemtl plain
f 1//1 2//1 3//1
f -5/-4/1 -4/-3/1 -3/-2/1 -2/-1/1 -1/1/1
usemtl red
f 1 3 5
"""


@pytest.mark.skipif(not _have_reference(), reason="glm / stb come from the reference's third_party by include path")
def test_model_obj_import(tmp_path):
    """Model<HasTexCoord> (model.h:14-114 without assimp): fan triangulation of polygons, v / v/vt / v//vn /
    v/vt/vn corners, negative indices, one mesh per material in order of first use (texture slot = mesh index),
    host-side transform, the 60- / 36-byte Face layouts, and the diffuse texture looked up at
    <model dir>/../textures/<basename of map_Kd> (model.h:80-88) and expanded to RGBA8."""
    subprocess.run(["make", "-s", "-C", API, "build/libray_tracing.a"], check=True)
    exe = str(tmp_path / "obj_dump")
    ref = "/root/reference/third_party"
    subprocess.run(["nvcc", "-std=c++17", "-O1", "-gencode", "arch=compute_100a,code=sm_100a", "--expt-relaxed-constexpr",
                    "--expt-extended-lambda", "-w", "-I" + API, f"-I{ref}/glm", f"-I{ref}/stb", "-o", exe,
                    os.path.join(ROOT, "tests", "obj_dump.cu"), os.path.join(BUILD, "libray_tracing.a"),
                    "-L" + os.path.dirname(API), "-lrtcore", "-Xlinker", "-rpath", "-Xlinker", os.path.dirname(API)], check=True)
    (tmp_path / "r" / "models").mkdir(parents=True)
    (tmp_path / "r" / "textures").mkdir()
    (tmp_path / "r" / "models" / "a.obj").write_text(_OBJ)
    (tmp_path / "r" / "models" / "a.mtl").write_text("newmtl red\nKd 1 0 0\nmap_Kd some/dir/tex.ppm\nnewmtl plain\nKd 1 1 1\n")
    (tmp_path / "r" / "textures" / "tex.ppm").write_bytes(b"P6\n2 1\n255\n" + bytes([1, 2, 3, 10, 11, 12]))
    out = subprocess.run([exe, str(tmp_path / "r" / "models" / "a.obj")], capture_output=True, text=True, check=True).stdout
    lines = out.strip().split("\n")

    v = np.array([[0, 0, 0], [1, 0, 0], [1, 1, 0], [0, 1, 0], [0.5, 1.5, 0.25]], np.float32) * np.float32(2) + \
        np.array([1, 2, 3], np.float32)  # the tool's transform: translate(1,2,3) * scale(2); exact in float
    vt = np.array([[0, 0], [1, 0], [1, 1], [0, 1]], np.float32)
    z = None  # no texcoord on the corner -> (0, 0)
    red = [((0, 0), (1, 1), (2, 2)), ((0, 0), (2, 2), (3, 3)), ((0, z), (2, z), (4, z))]
    plain = [((0, z), (1, z), (2, z)), ((0, 0), (1, 1), (2, 2)), ((0, 0), (2, 2), (3, 3)), ((0, 0), (3, 3), (4, 0))]

    def expect(faces, with_uv):
        rows = []
        for f in faces:
            for vi, ti in f:
                uv = vt[ti] if ti is not None else np.zeros(2, np.float32)
                rows.append(list(v[vi]) + (list(uv) if with_uv else []))
        return np.array(rows, np.float32)

    assert lines[0].split() == ["2", "2"]
    at = 1
    for with_uv in (True, False):
        for faces, tex_id in ((red, 0), (plain, -1)):
            assert lines[at].split() == ["mesh", str(len(faces)), str(tex_id)], lines[at]
            got = np.array([[float.fromhex(x) for x in l.split()] for l in lines[at + 1: at + 1 + 3 * len(faces)]], np.float32)
            assert np.array_equal(got, expect(faces, with_uv)), (with_uv, tex_id)
            at += 1 + 3 * len(faces)
    assert lines[at].split() == ["tex", "1", "2", "8", "1", "2", "3", "255", "10", "11", "12", "255"], lines[at]
    assert lines[at + 1].split() == ["tex", "0", "0", "0"]
    assert lines[at + 2].split() == ["sizes", "60", "36"]  # bvh.cuh:35-65: Face<true> / Face<false>
