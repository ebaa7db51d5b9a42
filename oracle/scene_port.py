"""TEST INFRASTRUCTURE (oracle/): plain-Python restatement of the reference scene loader.

Follows scene_reader.py:11-46 (tokeniser / comment stripping), :62-104 (`.obj`: `v`, `f`, negative
indices, fan triangulation), :139-188 (`.sdl` keywords), :5-8 (`calc_normal`) and
vector.py:143-146, 109-111, 164-165, 172-173 (cross product, size, area, normalise) in Python
floats, i.e. the same IEEE operations as the reference, and flattens the result into the arrays the
C oracle (pt_oracle.c `pto_scene`) consumes.  Pinned by tests/test_oracle_golden.py against arrays
dumped from the reference's own `Scene` object (tests/golden/cornell_scene.npz).
"""
import os
from math import sqrt

import numpy as np


def _lines(path):
    # scene_reader.py:36-46
    out = []
    with open(path, "r") as f:
        for line in f.readlines():
            i = 0
            while line[i] == " ":
                i += 1
            line = line[i:]
            if line[0] != "#":
                if "#" in line:
                    line = line.split("#")[0]
                out.append(line.replace("\n", "").replace("\t", " "))
    return out


def _tokens(line):
    return [t for t in line.split(" ") if t not in ("", " ")]   # scene_reader.py:16


def _cross(u, v):
    return (u[1] * v[2] - u[2] * v[1], u[2] * v[0] - u[0] * v[2], u[0] * v[1] - u[1] * v[0])


def _size(v):
    return sqrt(sum([c ** 2 for c in v]))


def _sub(a, b):
    return tuple(-y + x for x, y in zip(a, b))   # V.__sub__ = self + (-vec)  (vector.py:84-86, 63-67)


def load_obj(path):
    """-> (triangles [n][3][3], normals [n][3], areas [n]) as Python-float lists."""
    verts, tris, normals, areas = [], [], [], []
    for line in _lines(path):
        tok = _tokens(line)
        if not tok:
            continue
        if tok[0] == "v":
            verts.append(tuple(float(x) for x in tok[1:]))
        elif tok[0] == "f":
            face = []
            for i in tok[1:]:
                i = int(i)
                face.append(len(verts) + i if i < 0 else i - 1)
            fs = [(face[0], face[i], face[i + 1]) for i in range(1, len(face) - 1)] if len(face) > 3 else [face]
            for f in fs:
                t = (verts[f[0]][:3], verts[f[1]][:3], verts[f[2]][:3])
                c = _cross(_sub(t[1], t[0]), _sub(t[2], t[0]))
                inv = 1 / _size(c)
                tris.append(t)
                normals.append(tuple(x * inv for x in c))
                areas.append(_size(c) / 2)
    return tris, normals, areas


def load_scene(path, width=None, height=None):
    """-> dict of numpy arrays / scalars with the pto_scene layout."""
    sc = {"objects": [], "npaths": None, "seed": None, "tonemapping": None, "background": None}
    light = None
    for line in _lines(path):
        tok = _tokens(line)
        if not tok:
            continue
        k = tok[0]
        if k == "eye":
            sc["eye"] = [float(t) for t in tok[1:4]]
        elif k == "size":
            sc["width"], sc["height"] = int(tok[1]), int(tok[2])
        elif k == "ortho":
            sc["ortho"] = [float(t) for t in tok[1:5]]
        elif k == "background":
            sc["background"] = [float(t) for t in tok[1:4]]
        elif k == "ambient":
            sc["ambient"] = float(tok[1])
        elif k == "light":
            light = load_obj(os.path.join(os.path.dirname(path), tok[1]))
            sc["light_color"] = [float(t) for t in tok[2:6]]
        elif k == "npaths":
            sc["npaths"] = int(tok[1])
        elif k == "tonemapping":
            sc["tonemapping"] = float(tok[1])
        elif k == "seed":
            sc["seed"] = int(tok[1])
        elif k == "object":
            geo = load_obj(os.path.join(os.path.dirname(path), tok[1]))
            sc["objects"].append((geo, [float(t) for t in tok[2:10]]))
        elif k == "output":
            sc["output"] = os.path.join(os.path.dirname(path), tok[1])
    if width is not None:
        sc["width"] = int(width)
    if height is not None:
        sc["height"] = int(height)
    verts, normals, tri_obj = [], [], []
    for oi, (geo, _) in enumerate(sc["objects"]):
        verts += geo[0]
        normals += geo[1]
        tri_obj += [oi] * len(geo[0])
    n_tri_obj = len(verts)
    verts += light[0]
    normals += light[1]
    tri_obj += [len(sc["objects"])] * len(light[0])
    flat = {
        "verts": np.asarray(verts, np.float64).reshape(-1, 3, 3),
        "normals": np.asarray(normals, np.float64).reshape(-1, 3),
        "tri_obj": np.asarray(tri_obj, np.int32),
        "mat": np.asarray([m for _, m in sc["objects"]], np.float64).reshape(-1, 8),
        "light_areas": np.asarray(light[2], np.float64),
        "n_tri_obj": n_tri_obj,
        "eye": np.asarray(sc["eye"], np.float64),
        "ortho": np.asarray(sc["ortho"], np.float64),
        "width": sc["width"], "height": sc["height"],
        "ambient": sc["ambient"],
        "light_color": np.asarray(sc["light_color"][:3], np.float64),
        "npaths": sc["npaths"], "seed": sc["seed"],
    }
    return flat
