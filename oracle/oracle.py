"""TEST INFRASTRUCTURE (oracle/): ctypes front-end of the FP64 C restatement (pt_oracle.c).

Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / `--impl reference` legs may
import this.  Nothing under pathtracerpython_b200/ does.
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "_build", "libpt_oracle.so")
_lib = None


def build(force=False):
    src = os.path.join(_HERE, "pt_oracle.c")
    if force or not os.path.exists(_SO) or os.path.getmtime(_SO) < os.path.getmtime(src):
        subprocess.check_call(["make", "-C", _HERE, "-s"])
    return _SO


class _Scene(C.Structure):
    _fields_ = [("n_tri", C.c_int32), ("n_tri_obj", C.c_int32), ("n_obj", C.c_int32), ("n_light_tri", C.c_int32),
                ("verts", C.c_void_p), ("normals", C.c_void_p), ("tri_obj", C.c_void_p), ("mat", C.c_void_p),
                ("light_areas", C.c_void_p), ("eye", C.c_double * 3), ("ortho", C.c_double * 4),
                ("width", C.c_int32), ("height", C.c_int32), ("ambient", C.c_double),
                ("light_color", C.c_double * 3)]


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = C.CDLL(_SO)
        _lib.pto_make_image.restype = C.c_int
        _lib.pto_intersect_one.restype = C.c_int
        _lib.pto_num_threads.restype = C.c_int
        _lib.pto_set_threads.restype = C.c_int
        _lib.pto_set_threads.argtypes = [C.c_int]
    return _lib


def _p(a):
    return None if a is None else a.ctypes.data_as(C.c_void_p)


class Oracle:
    """Holds a flattened scene (dict from oracle.scene_port.load_scene or any dict with the same
    keys) and runs the C restatement on it."""

    def __init__(self, flat, width=None, height=None):
        self.flat = flat
        self._keep = {k: np.ascontiguousarray(flat[k], dt) for k, dt in
                      [("verts", np.float64), ("normals", np.float64), ("tri_obj", np.int32),
                       ("mat", np.float64), ("light_areas", np.float64)]}
        s = _Scene()
        s.n_tri = self._keep["verts"].shape[0]
        s.n_tri_obj = int(flat["n_tri_obj"])
        s.n_obj = self._keep["mat"].shape[0]
        s.n_light_tri = self._keep["light_areas"].shape[0]
        for k in ("verts", "normals", "tri_obj", "mat", "light_areas"):
            setattr(s, k, self._keep[k].ctypes.data)
        s.eye[:] = [float(x) for x in flat["eye"]]
        s.ortho[:] = [float(x) for x in flat["ortho"]]
        s.width = int(width if width is not None else flat["width"])
        s.height = int(height if height is not None else flat["height"])
        s.ambient = float(flat["ambient"])
        s.light_color[:] = [float(x) for x in flat["light_color"][:3]]
        self.s = s
        self.width, self.height = s.width, s.height

    def primary(self, pixels=None):
        n = self.width * self.height if pixels is None else len(pixels)
        px = None if pixels is None else np.ascontiguousarray(pixels, np.int64)
        tri = np.empty(n, np.int32)
        d2 = np.empty(n, np.float64)
        pt = np.empty((n, 3), np.float64)
        lib().pto_primary(C.byref(self.s), _p(px), C.c_int64(n), _p(tri), _p(d2), _p(pt))
        return tri, d2, pt

    def closest(self, rays):
        rays = np.ascontiguousarray(rays, np.float64).reshape(-1, 6)
        n = rays.shape[0]
        tri = np.empty(n, np.int32)
        d2 = np.empty(n, np.float64)
        pt = np.empty((n, 3), np.float64)
        lib().pto_closest(C.byref(self.s), _p(rays), C.c_int64(n), _p(tri), _p(d2), _p(pt))
        return tri, d2, pt

    def trace(self, spp, bounces, seed, pixels=None, details=True):
        n = self.width * self.height if pixels is None else len(pixels)
        px = None if pixels is None else np.ascontiguousarray(pixels, np.int64)
        out = {"radiance": np.zeros((n, 3), np.float64)}
        if details:
            sh = (n, spp, bounces)
            out.update(tri=np.empty(sh, np.int32), point=np.empty(sh + (3,), np.float64),
                       contrib=np.empty(sh + (3,), np.float64), k_after=np.empty(sh, np.float64),
                       lobe=np.empty(sh, np.int8), occ=np.empty(sh, np.int8), leak=np.empty(sh, np.int8))
        rc = np.zeros(2, np.int64)
        g = out.get
        lib().pto_trace(C.byref(self.s), _p(px), C.c_int64(n), C.c_int32(spp), C.c_int32(bounces), C.c_uint32(seed),
                        _p(g("tri")), _p(g("point")), _p(g("contrib")), _p(g("k_after")), _p(g("lobe")),
                        _p(g("occ")), _p(g("leak")), _p(out["radiance"]), _p(rc))
        out["closest_rays"], out["shadow_rays"] = int(rc[0]), int(rc[1])
        return out

    # ---- known answers (tests/golden/kat_misc.npz)
    def light_sample(self, seed, positions):
        """positions [n,4] = (pixel, sample, bounce, light sample index) -> (pick [n], point [n,3])"""
        pos = np.ascontiguousarray(positions, np.int32).reshape(-1, 4)
        pick = np.empty(len(pos), np.int32)
        pt = np.empty((len(pos), 3), np.float64)
        lib().pto_light_sample(C.byref(self.s), C.c_uint32(seed), _p(pos), C.c_int64(len(pos)), _p(pick), _p(pt))
        return pick, pt

    def primary_dirs(self, rays):
        r = np.ascontiguousarray(rays, np.int64)
        out = np.empty((len(r), 3), np.float64)
        lib().pto_primary_dirs(C.byref(self.s), _p(r), C.c_int64(len(r)), _p(out))
        return out

    def make_image(self, radiance):
        radiance = np.ascontiguousarray(radiance, np.float64).reshape(-1, 3)
        img = np.empty((self.height, self.width, 3), np.uint8)
        rc = lib().pto_make_image(_p(radiance), C.c_int32(self.width), C.c_int32(self.height), _p(img))
        if rc != 0:
            raise IndexError("make_image index mapping leaves the image (reference: utils.py:153-157)")
        return img


def intersect_one(pr, vr, tri):
    pr = np.ascontiguousarray(pr, np.float64)
    vr = np.ascontiguousarray(vr, np.float64)
    tri = np.ascontiguousarray(tri, np.float64).reshape(9)
    P = np.zeros(3, np.float64)
    hit = lib().pto_intersect_one(_p(pr), _p(vr), _p(tri), _p(P))
    return (P if hit else None)


def rotate_y(angles, vecs):
    a = np.ascontiguousarray(angles, np.float64)
    v = np.ascontiguousarray(vecs, np.float64).reshape(-1, 3)
    out = np.empty_like(v)
    lib().pto_rotate_y(_p(a), _p(v), C.c_int64(len(a)), _p(out))
    return out


def philox(ctr, key):
    c = np.asarray(ctr, np.uint32)
    k = np.asarray(key, np.uint32)
    o = np.zeros(4, np.uint32)
    lib().pto_philox(_p(c), _p(k), _p(o))
    return o


def num_threads():
    return lib().pto_num_threads()


def host_cores():
    """Cores this process may run on (cgroup / affinity aware)."""
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def set_threads(n=0):
    """OpenMP threads of the C restatement; n <= 0: every core this process may use, whatever OMP_NUM_THREADS says
    (torchrun exports OMP_NUM_THREADS=1 to its children).  Returns the count in force."""
    return lib().pto_set_threads(int(n) if n > 0 else host_cores())
