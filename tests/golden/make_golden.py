#!/usr/bin/env python3
"""Generates the committed golden vectors by RUNNING the unmodified reference
(/root/reference, import stubs in oracle/stubs) through oracle/ref_driver.py.

Run in the build container only (the reference is not on the GPU box):
    python tests/golden/make_golden.py
Outputs (all under tests/golden/):
    cornell_scene.npz      flattened reference `Scene` of objs/cornellroom.sdl
    kat_intersect.npz      utils.intersect known answers (hits, misses, behind-origin, edge, parallel)
    kat_misc.npz           rotate / pick_random_triangle / sample_random_pt / make_image / linspace rays
    cornell_primary.npz    primary closest-hit triangle ids: full 40x40, crops of 512^2 and 1024^2
    cornell_paths_40.npz   40x40, 2 spp, 3 bounces, seed 9: per (pixel,sample,bounce) records + image
"""
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
from oracle import ref_driver as rd  # noqa: E402
from oracle.philox import Stream  # noqa: E402

OUT = os.path.dirname(os.path.abspath(__file__))
SDL = os.path.join(rd.REF_ROOT, "objs", "cornellroom.sdl")
SEED = 9  # cornellroom.sdl:19


def dump_scene():
    sc = rd.load_scene(SDL)
    verts, normals, tri_obj, mats = [], [], [], []
    for oi, o in enumerate(sc.objects):
        g = o["geometry"]
        verts += [[list(v.coords) for v in t] for t in g.triangles]
        normals += [list(n.coords) for n in g.normals]
        tri_obj += [oi] * len(g.triangles)
        mats.append([o[k] for k in ("red", "green", "blue", "ka", "kd", "ks", "kt", "n")])
    n_tri_obj = len(verts)
    g = sc.light_obj
    verts += [[list(v.coords) for v in t] for t in g.triangles]
    normals += [list(n.coords) for n in g.normals]
    tri_obj += [len(sc.objects)] * len(g.triangles)
    np.savez_compressed(
        os.path.join(OUT, "cornell_scene.npz"),
        verts=np.array(verts), normals=np.array(normals), tri_obj=np.array(tri_obj, np.int32),
        mat=np.array(mats), light_areas=np.array(g.areas), n_tri_obj=n_tri_obj,
        eye=np.array(sc.eye), ortho=np.array(sc.ortho), width=sc.width, height=sc.height,
        ambient=sc.ambient, light_color=np.array(sc.light_color), npaths=sc.npaths, seed=sc.seed,
        tonemapping=sc.tonemapping, background=np.array(sc.background))
    return sc


def kat_intersect(sc):
    _, ru, _ = rd.load_reference()
    rng = np.random.default_rng(1)
    tris = [t for o in sc.objects for t in o["geometry"].triangles] + list(sc.light_obj.triangles)
    rays, tri_ix, hit, pts = [], [], [], []

    def run(o, d, ti):
        with np.errstate(all="ignore"):
            try:
                p = ru.intersect((np.array(o, float), np.array(d, float)), tris[ti])
                hit.append(1)
                pts.append(p)
            except ru.NoIntersection:
                hit.append(0)
                pts.append(np.zeros(3))
        rays.append(list(o) + list(d))
        tri_ix.append(ti)

    for _ in range(600):   # rays from inside the room aimed at a random point of a random triangle's plane
        ti = int(rng.integers(len(tris)))
        t = np.array([list(v.coords) for v in tris[ti]])
        w = rng.uniform(-0.3, 1.3, 3)
        w /= w.sum()
        target = w @ t
        o = np.array([rng.uniform(-3.5, 3.5), rng.uniform(-3.5, 3.5), rng.uniform(-32, 5)])
        d = (target - o) * rng.uniform(0.2, 3.0)
        if rng.random() < 0.3:
            d = -d                                    # triangle BEHIND the origin still hits (A-Q1)
        run(o, d, ti)
    # exact edge / vertex targets (NaN in in_triangle -> miss, A-Q15) and parallel rays
    for ti in range(0, len(tris), 3):
        t = np.array([list(v.coords) for v in tris[ti]])
        o = np.array([0.1, 0.2, 5.7])
        run(o, t[0] - o, ti)
        run(o, (t[0] + t[1]) / 2 - o, ti)
        run(o, t[1] - t[0], ti)                       # parallel to the plane
    np.savez_compressed(os.path.join(OUT, "kat_intersect.npz"), rays=np.array(rays),
                        tri=np.array(tri_ix, np.int32), hit=np.array(hit, np.int8), point=np.array(pts))


def kat_misc(sc):
    rm, ru, _ = rd.load_reference()
    rng = np.random.default_rng(2)
    # rotate about y (main.py:148-162)
    ang = rng.uniform(0, np.pi, 64)
    vec = rng.normal(size=(64, 3))
    rot = np.array([rm.rotate(np.array((0, 1, 0)), a, v) for a, v in zip(ang, vec)])
    # light sampling with the injected stream (utils.py:21-46)
    st = Stream(SEED)
    ru.uniform = st.uniform
    picks, lpts = [], []
    for pix in range(256):
        st.at(pix, 3, 1, 0)
        i = ru.pick_random_triangle(sc.light_obj.areas)
        p = ru.sample_random_pt(sc.light_obj.triangles[i])
        picks.append(i)
        lpts.append(p)
    # make_image (utils.py:150-161) on signed radiance, 24x24
    rad = rng.normal(0.3, 0.4, size=(24 * 24, 3))
    with np.errstate(all="ignore"):
        img = np.asarray(ru.make_image(-1, -1, 1, 1, 24, 24, [(c, None) for c in rad]))
    # primary rays (utils.py:55-69) at 7x7 over the shipped window
    pts = ru.make_screen_pts(*sc.ortho, 7, 7)
    rays = ru.make_rays(sc.eye, pts)
    np.savez_compressed(os.path.join(OUT, "kat_misc.npz"), rot_angle=ang, rot_in=vec, rot_out=rot,
                        light_pick=np.array(picks, np.int32), light_pt=np.array(lpts),
                        img_radiance=rad, img_u8=img, ray_dirs=np.array([r[1] for r in rays]))


def primary(sc):
    out = {}
    t = time.time()
    full = rd.trace_parallel(SDL, None, None, 1, 1, SEED, with_shadow_bits=False)
    out["tri_40"] = full["tri"][:, 0, 0]
    rng = np.random.default_rng(3)
    for res in (512, 1024):
        px = np.sort(rng.choice(res * res, 3000, replace=False))
        r = rd.trace_parallel(SDL, res, res, 1, 1, SEED, pixels=px, with_shadow_bits=False)
        out[f"pixels_{res}"] = px
        out[f"tri_{res}"] = r["tri"][:, 0, 0]
        out[f"point_{res}"] = r["point"][:, 0, 0]
    print("primary goldens", time.time() - t, "s")
    np.savez_compressed(os.path.join(OUT, "cornell_primary.npz"), **out)


def paths(sc):
    t = time.time()
    r = rd.trace_parallel(SDL, None, None, 2, 3, SEED)
    img = rd.reference_image(sc, r["radiance"])
    print("path goldens", time.time() - t, "s")
    np.savez_compressed(os.path.join(OUT, "cornell_paths_40.npz"), spp=2, bounces=3, seed=SEED,
                        tri=r["tri"], point=r["point"], contrib=r["contrib"], k_after=r["k_after"],
                        lobe=r["lobe"], occ=r["occ"], leak=r["leak"], radiance=r["radiance"], image=img)


def obj_corners():
    """The reference's own .obj parser (scene_reader.Obj, unmodified) on every corner case of tests/obj_corners.py:
    what it returns - vertices as read (any number of coordinates), faces as stored, triangles / normals / areas - or
    the class of the exception it raises, and how many 'Skipping command' reports it prints."""
    import contextlib
    import io
    import json
    import tempfile
    sys.path.insert(0, os.path.dirname(OUT))
    from obj_corners import CASES
    _, _, rs = rd.load_reference()
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for name, text in CASES.items():
            path = os.path.join(tmp, name + ".obj")
            with open(path, "w", newline="") as f:
                f.write(text)
            buf = io.StringIO()
            try:
                with contextlib.redirect_stdout(buf), np.errstate(all="ignore"):
                    o = rs.Obj(path)
            except Exception as e:      # noqa: BLE001 - the class is the golden
                out[name] = {"raises": type(e).__name__}
                continue
            out[name] = {"raises": None, "skipped": buf.getvalue().count("Skipping command"),
                         "vertexes": [[float(c) for c in v.coords] for v in o.vertexes],
                         "faces": [[int(i) for i in f] for f in o.faces],
                         "triangles": [[[float(c) for c in list(v.coords)[:3]] for v in t] for t in o.triangles],
                         "normals": [[float(c) for c in n.coords] for n in o.normals],
                         "areas": [float(a) for a in o.areas]}
    with open(os.path.join(OUT, "obj_corners.json"), "w") as f:
        json.dump(out, f, indent=0)
    print("obj corner goldens:", {k: v["raises"] or "ok" for k, v in out.items()})


if __name__ == "__main__":
    if "--obj-corners" in sys.argv:      # only the parser goldens (seconds); the rest takes minutes
        obj_corners()
        sys.exit(0)
    scene = dump_scene()
    obj_corners()
    kat_intersect(scene)
    kat_misc(scene)
    primary(scene)
    paths(scene)
    for f in sorted(os.listdir(OUT)):
        if f.endswith(".npz"):
            print(f, os.path.getsize(os.path.join(OUT, f)))
