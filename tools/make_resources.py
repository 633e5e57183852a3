"""Synthesises the two resource files the reference's scenes load but does not ship (.gitignore:1-2):
resources/bunny.obj (scenes/bunny.cu:102) and resources/earthmap.jpg (scenes/birthday.cu:80), SURVEY §8d.
usage: python tools/make_resources.py [dir]   (default: ./resources)"""
import importlib, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np


def main(out="resources", n_faces=69451, tex_hw=(1024, 2048)):
    rt = importlib.import_module("ray-tracing-cuda_b200")
    os.makedirs(out, exist_ok=True)
    rt.scenes.write_obj(os.path.join(out, "bunny.obj"), rt.scenes.bunny_mesh(n_faces))
    tex = rt.scenes.earth_texture(*tex_hw)
    rt.write_image(os.path.join(out, "earthmap.jpg"), np.ascontiguousarray(tex[..., :3]))
    return out


if __name__ == "__main__":
    print(main(*(sys.argv[1:2])))
