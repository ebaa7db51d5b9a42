#!/usr/bin/env python3
"""Writes the Cornell-box fixture (pathtracerpython_b200/scenes/cornell/) from the golden scene
dump (tests/golden/cornell_scene.npz, itself produced from the reference's `Scene` object by
make_golden.py).  The files use the reference's `.sdl` / `.obj` grammar with
repr-exact decimals, one `f` line per triangle, objects in the reference's order, so parsing them
reproduces the golden arrays bit for bit (tests/test_scene.py)."""
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
OUT = os.path.join(os.path.dirname(os.path.dirname(HERE)), "pathtracerpython_b200", "scenes", "cornell")
NAMES = ["wall_left", "wall_right", "floor", "wall_back", "ceiling", "block_small", "block_tall"]


def fmt(x):
    return repr(float(x))


def write_obj(path, tris, title):
    verts, index, faces = [], {}, []
    for t in tris:
        f = []
        for v in t:
            key = tuple(v.tolist())
            if key not in index:
                index[key] = len(verts) + 1
                verts.append(key)
            f.append(index[key])
        faces.append(f)
    with open(path, "w") as fh:
        fh.write(f"# {title}: {len(verts)} vertices, {len(faces)} triangles\n")
        for v in verts:
            fh.write("v " + " ".join(fmt(c) for c in v) + "\n")
        for f in faces:
            fh.write("f " + " ".join(str(i) for i in f) + "\n")


def main():
    g = np.load(os.path.join(HERE, "cornell_scene.npz"))
    os.makedirs(OUT, exist_ok=True)
    verts, tri_obj, mat = g["verts"], g["tri_obj"], g["mat"]
    n_obj = mat.shape[0]
    write_obj(os.path.join(OUT, "light_panel.obj"), verts[tri_obj == n_obj], "ceiling light panel")
    for oi in range(n_obj):
        write_obj(os.path.join(OUT, NAMES[oi] + ".obj"), verts[tri_obj == oi], NAMES[oi])
    with open(os.path.join(OUT, "cornellroom.sdl"), "w") as fh:
        fh.write("# Cornell box fixture: same camera, light, objects and coefficients as the scene the\n"
                 "# reference renderer ships with; written by tests/golden/make_scene_files.py\n")
        fh.write("eye " + " ".join(fmt(c) for c in g["eye"]) + "\n")
        fh.write(f"size {int(g['width'])} {int(g['height'])}\n")
        fh.write("ortho " + " ".join(fmt(c) for c in g["ortho"]) + "\n")
        fh.write("background " + " ".join(fmt(c) for c in g["background"]) + "\n")
        fh.write(f"ambient {fmt(g['ambient'])}\n")
        fh.write("light light_panel.obj " + " ".join(fmt(c) for c in g["light_color"]) + "\n")
        fh.write(f"npaths {int(g['npaths'])}\ntonemapping {fmt(g['tonemapping'])}\nseed {int(g['seed'])}\n")
        fh.write("# object <file.obj> red green blue ka kd ks kt n\n")
        for oi in range(n_obj):
            fh.write(f"object {NAMES[oi]}.obj " + " ".join(fmt(c) for c in mat[oi]) + "\n")
        fh.write("output cornell.pnm\n")


if __name__ == "__main__":
    main()
