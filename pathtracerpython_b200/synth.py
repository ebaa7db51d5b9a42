"""Synthetic tessellated scenes (BASELINE.json configs 4 and 5): every triangle of a scene is split
into n x n coplanar triangles, per object, so materials, the light and the object order are unchanged.
Cornell box: 32 * n^2 triangles (n = 177 -> 1 002 528, n = 559 -> 9 999 392).  Deterministic, no RNG.
`write_scene` stores the result with the reference's `.sdl` / `.obj` grammar and repr-exact decimals.
"""
import os

import numpy as np

from . import _lib
from .scene import Obj, Scene


def tessellate_triangles(tris, n):
    """tris [m,3,3] float64 -> [m*n*n,3,3]; same winding, row-major over the barycentric grid."""
    tris = np.asarray(tris, np.float64)
    m = tris.shape[0]
    A, B, C = tris[:, 0], tris[:, 1], tris[:, 2]
    e1, e2 = (B - A), (C - A)

    def pt(i, j):          # [k] grid indices -> [m,k,3]
        return A[:, None, :] + (i / n)[None, :, None] * e1[:, None, :] + (j / n)[None, :, None] * e2[:, None, :]

    ii, jj = np.meshgrid(np.arange(n), np.arange(n), indexing="ij")
    up = (ii + jj) <= n - 1
    iu, ju = ii[up].astype(np.float64), jj[up].astype(np.float64)          # upright: (i,j) (i+1,j) (i,j+1)
    dn = (ii + jj) <= n - 2
    idn, jdn = ii[dn].astype(np.float64), jj[dn].astype(np.float64)        # inverted: (i+1,j) (i+1,j+1) (i,j+1)
    upt = np.stack([pt(iu, ju), pt(iu + 1, ju), pt(iu, ju + 1)], axis=2)           # [m,ku,3,3]
    dnt = np.stack([pt(idn + 1, jdn), pt(idn + 1, jdn + 1), pt(idn, jdn + 1)], axis=2)
    out = np.concatenate([upt, dnt], axis=1)                                       # [m, n*n, 3, 3]
    return np.ascontiguousarray(out.reshape(m * n * n, 3, 3))


def _obj_from_triangles(tris, path):
    verts = tris.reshape(-1, 3)
    faces = np.arange(verts.shape[0], dtype=np.int64).reshape(-1, 3)
    return Obj.from_arrays(verts, faces, path)


def tessellate(scene, n):
    """Scene -> new Scene whose meshes are the n x n subdivision of the input's."""
    out = Scene()
    for k in ("eye", "width", "height", "ortho", "background", "ambient", "light_color", "npaths", "tonemapping", "seed"):
        setattr(out, k, getattr(scene, k))
    out.light_obj = _obj_from_triangles(tessellate_triangles(scene.light_obj.triangles, n), "light")
    for o in scene.objects:
        d = dict(o)
        d["geometry"] = _obj_from_triangles(tessellate_triangles(o["geometry"].triangles, n), "object")
        out.objects.append(d)
    return out


def warp(scene, amplitude=0.04, freq=2.3):
    """Scene -> new Scene with every vertex moved by a smooth vector field of its position, p + a * (sin(f p.y + 1),
    sin(f p.z + 2), sin(f p.x + 3)): shared vertices move together (the meshes stay watertight, the light stays in front
    of the ceiling), but no two triangles of a tessellated wall lie in one plane any more.  A tessellated + warped
    Cornell box is the worst case for the LBVH path's supporting-plane culling (10^6 distinct planes: the scene goes
    without it, ptb_scene_plane_culling == 0) where the flat tessellation (32 planes) is its best case."""
    def field(v):
        v = np.asarray(v, np.float64)
        d = np.stack([np.sin(freq * v[..., 1] + 1.0), np.sin(freq * v[..., 2] + 2.0), np.sin(freq * v[..., 0] + 3.0)], -1)
        return v + amplitude * d

    out = Scene()
    for k in ("eye", "width", "height", "ortho", "background", "ambient", "light_color", "npaths", "tonemapping", "seed"):
        setattr(out, k, getattr(scene, k))
    out.light_obj = _obj_from_triangles(field(scene.light_obj.triangles), "light")
    for o in scene.objects:
        d = dict(o)
        d["geometry"] = _obj_from_triangles(field(o["geometry"].triangles), "object")
        out.objects.append(d)
    return out


def write_obj(path, obj):
    """One `v` line per triangle corner and one `f` line per triangle, decimals that read back bit-exactly.  Meshes of
    more than a few thousand triangles go through the library's writer (ptb_obj_write) when it is built."""
    n = len(obj.triangles)
    if n > 4096 and os.path.exists(_lib.LIB_PATH):
        _lib.obj_write(path, obj.triangles.reshape(-1, 3), np.arange(3 * n, dtype=np.int64).reshape(-1, 3), f"{n} triangles")
        return
    with open(path, "w") as f:
        f.write(f"# {n} triangles\n")
        v = obj.triangles.reshape(-1, 3)
        f.write("".join(f"v {repr(float(a))} {repr(float(b))} {repr(float(c))}\n" for a, b, c in v))
        f.write("".join(f"f {3 * i + 1} {3 * i + 2} {3 * i + 3}\n" for i in range(n)))


def write_scene(scene, directory, name="scene.sdl"):
    """Write `scene` as <directory>/<name> plus one .obj per mesh; returns the .sdl path."""
    os.makedirs(directory, exist_ok=True)
    write_obj(os.path.join(directory, "light.obj"), scene.light_obj)
    lines = [f"eye {' '.join(repr(float(x)) for x in scene.eye)}", f"size {scene.width} {scene.height}",
             f"ortho {' '.join(repr(float(x)) for x in scene.ortho)}"]
    if scene.background is not None:
        lines.append(f"background {' '.join(repr(float(x)) for x in scene.background)}")
    lines.append(f"ambient {repr(float(scene.ambient))}")
    lines.append(f"light light.obj {' '.join(repr(float(x)) for x in scene.light_color)}")
    if scene.npaths is not None:
        lines.append(f"npaths {scene.npaths}")
    if scene.tonemapping is not None:
        lines.append(f"tonemapping {repr(float(scene.tonemapping))}")
    if scene.seed is not None:
        lines.append(f"seed {scene.seed}")
    for i, o in enumerate(scene.objects):
        fn = f"object_{i}.obj"
        write_obj(os.path.join(directory, fn), o["geometry"])
        lines.append(f"object {fn} " + " ".join(repr(float(o[k])) for k in ("red", "green", "blue", "ka", "kd", "ks", "kt", "n")))
    path = os.path.join(directory, name)
    with open(path, "w") as f:
        f.write("\n".join(lines) + "\n")
    return path
