"""TEST INFRASTRUCTURE (oracle/): tier-1 oracle = the UNMODIFIED reference functions, driven
single-process with an injected Philox stream.

Runs only where /root/reference is mounted (this container; never on the GPU box).  It imports the
reference's own modules (main.py, utils.py, scene_reader.py) behind three non-arithmetic import
stubs (oracle/stubs: ipdb, colorama, pyqtgraph) and restates ONLY the control flow of
main.py:186-280, which is inline in `main()` and cannot be imported:

  * closest hit      -> reference `intersect_objects`            (main.py:83-122)   called as is
  * shading          -> reference `compute_color`                (main.py:142-145)  called as is
  * bounce step      -> restated from main.py:236-268 using reference `rotate` (main.py:148-162)
  * accumulation     -> restated from main.py:224-231, 269-280
  * image            -> reference `make_image`                   (utils.py:150-161) called as is

`main.uniform` and `utils.uniform` are replaced by oracle.philox.Stream.uniform so every draw is a
pure function of (seed, pixel, sample, bounce, dim) - see oracle/philox.py for the dimension map.

Used by tests/golden/make_golden.py to produce the committed golden vectors that pin the C
restatement (oracle/pt_oracle.c) and, through it, the CUDA path.
"""
import contextlib
import io
import math
import os
import sys

import numpy as np

from .philox import Stream

REF_ROOT = "/root/reference"
_STUBS = os.path.join(os.path.dirname(os.path.abspath(__file__)), "stubs")

_ref = None


def load_reference(ref_root=REF_ROOT):
    """Import the reference modules (unmodified) and return (main, utils, scene_reader)."""
    global _ref
    if _ref is not None:
        return _ref
    if not os.path.isdir(ref_root):
        raise RuntimeError(f"reference not mounted at {ref_root}; goldens are under tests/golden/")
    for p in (_STUBS, ref_root):
        if p not in sys.path:
            sys.path.insert(0, p)
    import main as ref_main  # noqa: E402  (reference main.py)
    import utils as ref_utils  # noqa: E402
    import scene_reader as ref_scene_reader  # noqa: E402
    _ref = (ref_main, ref_utils, ref_scene_reader)
    return _ref


def load_scene(path, width=None, height=None):
    """reference Scene(path) (scene_reader.py:107-121), optionally overriding `size`."""
    _, _, sr = load_reference()
    with contextlib.redirect_stdout(io.StringIO()):
        scene = sr.Scene(path)
    if width is not None:
        scene.width = int(width)
    if height is not None:
        scene.height = int(height)
    return scene


def triangle_index_table(scene):
    """id(normal V object) -> global triangle index, in intersect_objects order
    (objects in file order, light mesh last: main.py:91-96)."""
    table = {}
    g = 0
    for obj in scene.objects:
        for n in obj["geometry"].normals:
            table[id(n)] = g
            g += 1
    for n in scene.light_obj.normals:
        table[id(n)] = g
        g += 1
    return table


def _shadow_replay(ref_main, ref_utils, scene, point):
    """Replay of main.py:30-59 with the reference's own helpers, to expose the discrete outcome
    (per light sample: occluded?, and the object index the loop variable `obj` leaks, main.py:42,70).
    Consumes dims 0..11 exactly like compute_shadow_rays."""
    occ = 0
    leak = len(scene.objects) - 1
    areas = scene.light_obj.areas
    for s in range(3):
        i = ref_utils.pick_random_triangle(areas)
        light_pt = ref_utils.sample_random_pt(scene.light_obj.triangles[i])
        v = light_pt - point
        v = v / np.linalg.norm(v)
        light_d2 = ref_utils.squared_dist(point, light_pt)
        done = False
        leak = len(scene.objects) - 1
        for oi, obj in enumerate(scene.objects):
            for tri in obj["geometry"].triangles:
                try:
                    p = ref_utils.intersect((point, v), tri)
                    d2 = ref_utils.squared_dist(p, point)
                    if d2 < ref_main.ZERO:
                        continue
                    if d2 < light_d2:
                        done = True
                        break
                except ref_utils.NoIntersection:
                    pass
            if done:
                leak = oi
                break
        if done:
            occ |= 1 << s
    return occ, leak


def trace(scene, spp, bounces, seed, pixels=None, with_shadow_bits=True):
    """Path-trace `pixels` (ray indices r = ix*H + iy, utils.py:64-69; default all) for `spp`
    samples and `bounces` bounces.  Returns a dict of arrays indexed [pixel, sample, bounce]."""
    ref_main, ref_utils, _ = load_reference()
    stream = Stream(seed)
    ref_main.uniform = stream.uniform
    ref_utils.uniform = stream.uniform
    W, H = scene.width, scene.height
    with np.errstate(all="ignore"):
        screen_pts = ref_utils.make_screen_pts(*scene.ortho, W, H)
        all_rays = ref_utils.make_rays(scene.eye, screen_pts)
    if pixels is None:
        pixels = np.arange(W * H)
    pixels = np.asarray(pixels, dtype=np.int64)
    tri_of = triangle_index_table(scene)
    n_obj = len(scene.objects)
    P = len(pixels)
    out = {
        "pixels": pixels,
        "tri": np.full((P, spp, bounces), -2, np.int32),      # -1 miss, -2 path already dead
        "point": np.zeros((P, spp, bounces, 3)),
        "contrib": np.zeros((P, spp, bounces, 3)),             # new_color * accumulated_k
        "k_after": np.zeros((P, spp, bounces)),
        "lobe": np.full((P, spp, bounces), -1, np.int8),       # 0 diffuse, 1 specular
        "occ": np.zeros((P, spp, bounces), np.int8),           # bit s = light sample s occluded
        "leak": np.full((P, spp, bounces), -1, np.int8),       # leaked object index (main.py:70)
        "radiance": np.zeros((P, 3)),                          # mean over samples (main.py:274-280)
    }
    eye = scene.eye
    with np.errstate(all="ignore"):
        for ip, r in enumerate(pixels):
            r = int(r)
            total = 0
            for s in range(spp):
                ray = all_rays[r]
                k = np.ones(1)                                   # accumulated_k[i_ray] (main.py:190)
                color = np.array((0, 0, 0))                      # main.py:179
                for b in range(bounces):
                    res = ref_main.intersect_objects(ray, scene.objects, scene.light_obj)
                    if res is None:
                        if ray is not None:
                            out["tri"][ip, s, b] = -1
                        ray = None
                        continue
                    point, normal, obj, is_light = res
                    out["tri"][ip, s, b] = tri_of[id(normal)]
                    out["point"][ip, s, b] = point
                    if is_light:
                        new_color = np.array(scene.light_color)  # main.py:215
                    else:
                        stream.at(r, s, b, 0)
                        new_color = ref_main.compute_color(scene, obj, point, normal)
                        assert stream.dim == 12, stream.dim
                        if with_shadow_bits:
                            stream.at(r, s, b, 0)
                            occ, leak = _shadow_replay(ref_main, ref_utils, scene, point)
                            out["occ"][ip, s, b] = occ
                            out["leak"][ip, s, b] = leak
                    color = color + new_color * k[0]             # main.py:230-231
                    out["contrib"][ip, s, b] = new_color * k[0]
                    # ---- new ray: restated from main.py:236-268 ----
                    old_ray = ray
                    if not is_light:
                        stream.at(r, s, b, 12)
                        rnd = stream.uniform(0, obj["kd"] + obj["ks"])
                        if rnd <= obj["kd"]:
                            phi = np.arccos(math.sqrt(stream.uniform(0, 1)))
                            theta = ref_main.TAU * stream.uniform(0, 1)
                            v = np.array((np.sin(phi) * np.cos(theta),
                                          np.sin(phi) * np.sin(theta),
                                          np.cos(phi)))
                            v = v / np.linalg.norm(v)
                            v = ref_main.rotate(np.array((0, 1, 0)), np.arccos(
                                np.dot(np.array((0, 1, 0)), normal)), v)
                            ray = (point, v)
                            k[0] *= obj["kd"] * np.dot(v, normal)
                            out["lobe"][ip, s, b] = 0
                        else:
                            _, old_v = old_ray
                            v = np.dot(normal, old_v) * 2 * normal - old_v
                            v = v / np.linalg.norm(v)
                            eye_v = eye - point
                            eye_v = eye_v / np.linalg.norm(eye_v)
                            v = ref_main.rotate(np.array((0, 1, 0)), np.arccos(
                                np.dot(np.array((0, 1, 0)), normal)), v)
                            ray = (point, v)
                            k[0] *= obj["ks"] * ((np.dot(eye_v, v)) ** obj["n"])
                            out["lobe"][ip, s, b] = 1
                    else:
                        ray = None
                    out["k_after"][ip, s, b] = k[0]
                total = total + color                            # main.py:269-271
            out["radiance"][ip] = total / spp                    # main.py:277
    out["n_obj"] = n_obj
    return out


def trace_parallel(scene_path, width, height, spp, bounces, seed, pixels=None, procs=None,
                   with_shadow_bits=True):
    """`trace` fanned out over processes by pixel chunks (pixels are independent: the stream is
    keyed by pixel).  Only a speed-up for golden generation; the arithmetic is `trace`."""
    import multiprocessing as mp
    scene = load_scene(scene_path, width, height)
    if pixels is None:
        pixels = np.arange(scene.width * scene.height)
    pixels = np.asarray(pixels, dtype=np.int64)
    procs = procs or os.cpu_count()
    chunks = [c for c in np.array_split(pixels, procs * 4) if len(c)]
    with mp.get_context("fork").Pool(procs) as pool:
        parts = pool.starmap(_trace_chunk, [(scene_path, width, height, spp, bounces, seed, c,
                                             with_shadow_bits) for c in chunks])
    out = {}
    for key in parts[0]:
        if key == "n_obj":
            out[key] = parts[0][key]
        else:
            out[key] = np.concatenate([p[key] for p in parts], axis=0)
    return out


def _trace_chunk(scene_path, width, height, spp, bounces, seed, pixels, with_shadow_bits):
    scene = load_scene(scene_path, width, height)
    return trace(scene, spp, bounces, seed, pixels, with_shadow_bits)


def reference_image(scene, radiance_full):
    """reference make_image (utils.py:150-161) on a full-frame radiance array [W*H, 3]."""
    _, ref_utils, _ = load_reference()
    inter = [(radiance_full[i], None) for i in range(radiance_full.shape[0])]
    with np.errstate(all="ignore"):
        im = ref_utils.make_image(*scene.ortho, scene.width, scene.height, inter)
    return np.asarray(im)
