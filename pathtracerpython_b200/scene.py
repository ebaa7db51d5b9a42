"""Scene input: the reference's `.sdl` / `.obj` grammar and its `Scene` / `Obj` interface
(scene_reader.py:49-188), re-implemented with packed numpy arrays and linear-time parsing.

Same surface as the reference: `Scene(path)` exposes eye, width, height, ortho, background, ambient,
light_obj, light_color, npaths, tonemapping, seed, objects (list of dicts with keys geometry, red,
green, blue, ka, kd, ks, kt, n) and output; `Obj` exposes triangles, normals, areas.  Same error
behaviour: unknown keywords are reported and skipped (scene_reader.py:102-104, 186-188), an unreadable
scene file prints 'Could not open file!' and re-raises OSError (:144-146), `i/j/k` face tokens raise
ValueError (:69).  Triangles/normals/areas are float64 arrays computed with the reference's operation
order (scene_reader.py:5-8; vector.py:109-111, 143-146, 164-165, 172-173), so they are bit-identical
to the reference's Python floats (tests/test_scene.py checks this against the golden dump).
"""
from os.path import dirname, join

import numpy as np


def _clean_lines(path):
    """scene_reader.py:29-46: strip leading blanks, drop `#` lines, cut trailing comments, tabs -> spaces."""
    with open(path, "r") as f:
        raw = f.readlines()
    out = []
    for line in raw:
        line = line.lstrip(" ")
        if not line:
            raise IndexError("string index out of range")      # reference: remove_spaces_from_start on '' (:31)
        if line[0] == "#":
            continue
        if "#" in line:
            line = line.split("#")[0]
        out.append(line.replace("\n", "").replace("\t", " "))
    return out


def _tokens(line):
    return [t for t in line.split(" ") if t]


class WrongDimensionError(Exception):
    """A face mixes vertices with different numbers of coordinates (the reference: vector.py:7-8, raised by
    V.__add__, vector.py:63-72, when calc_normal subtracts them, scene_reader.py:5-8)."""


def _native_parser_wanted(path, min_bytes):
    import os
    from . import _lib
    try:
        return os.path.getsize(path) >= min_bytes and os.path.exists(_lib.LIB_PATH)
    except OSError:
        return False                    # let the Python parser raise the reference's error


class Obj:
    """Triangle mesh (scene_reader.py:49-104).  triangles [n,3,3], normals [n,3], areas [n] float64."""

    NATIVE_MIN_BYTES = 1 << 16          # smaller files: the Python parser below is as quick as a library call

    def __init__(self, path, quiet=True, native=None):
        """native: None = ptb_obj_parse (libptb.so, multi-threaded) for files over NATIVE_MIN_BYTES when the
        library is built, True / False = force one parser.  Both give the same arrays and the same errors
        (tests/test_obj_native.py)."""
        if not quiet:
            print("Reading " + path)
        self.path = path
        if native is None:
            native = _native_parser_wanted(path, self.NATIVE_MIN_BYTES)
        tri_idx = None
        if native:
            from . import _lib
            try:
                self.vertexes, self.faces, skipped = _lib.obj_parse(path)
            except (ValueError, IndexError):
                # malformed or irregular input (a vertex that does not have 3 coordinates, a face index outside the
                # vertices read so far, ...): the line-by-line parser below raises exactly what the reference raises,
                # for the first offending line in file order
                native = False
            else:
                for line in skipped:
                    tok = _tokens(line.replace("\t", " "))
                    print(f"{path}\n\tSkipping command '{tok[0]}' ! \n\tParameters: {tok[1:]}")
        if not native:
            self.vertexes, self.faces, tri_idx = self._parse_python(path)
        self._finish(tri_idx, strict=True)

    @staticmethod
    def _parse_python(path):
        """The reference's parser line by line (scene_reader.py:62-104), its failure modes included: vertices keep
        however many coordinates their line has (vector.py:28-42) and a face is resolved against the vertices read SO
        FAR with Python's indexing (negative indices wrap, anything else outside raises IndexError); a face whose
        vertices differ in dimension raises WrongDimensionError (vector.py:63-72), one whose vertices have fewer than 3
        coordinates IndexError (cross_prod, vector.py:143-146), more than 3 are ignored.  Returns vertexes [n,3] (first
        three coordinates, missing ones 0), faces as the reference stores them and the resolved vertex index of every
        triangle corner."""
        verts, dims, faces, tri_idx = [], [], [], []
        for line in _clean_lines(path):
            tok = _tokens(line)
            if not tok:
                continue
            if tok[0] == "v":
                c = [float(x) for x in tok[1:]]
                if not c:
                    raise IndexError("tuple index out of range")            # V([]).coords[0], vector.py:37
                dims.append(len(c))
                verts.append((c + [0.0, 0.0])[:3])
            elif tok[0] == "f":
                nv = len(verts)
                face = []
                for t in tok[1:]:
                    i = int(t)                                   # 'i/j/k' -> ValueError, as the reference
                    face.append(nv + i if i < 0 else i - 1)
                if len(face) > 3:                                # fan about the first vertex (:75-78)
                    new = [(face[0], face[i], face[i + 1]) for i in range(1, len(face) - 1)]
                else:
                    new = [tuple(face)]
                for f in new:
                    if len(f) < 3:
                        raise IndexError("list index out of range")          # face[2], scene_reader.py:81-83
                    r = []
                    for k in f:
                        if not -nv <= k < nv:
                            raise IndexError("list index out of range")      # self.vertexes[face[k]]
                        r.append(k + nv if k < 0 else k)
                    if dims[r[1]] != dims[r[0]] or dims[r[2]] != dims[r[0]]:
                        raise WrongDimensionError(f"Cannot add vectors with dimensions {dims[r[1] if dims[r[1]] != dims[r[0]] else r[2]]} "
                                                  f"and {dims[r[0]]}")
                    if dims[r[0]] < 3:
                        raise IndexError("tuple index out of range")         # u[2] in cross_prod
                    faces.append(f)
                    tri_idx.append(r)
            else:
                print(f"{path}\n\tSkipping command '{tok[0]}' ! \n\tParameters: {tok[1:]}")
        return (np.asarray(verts, np.float64).reshape(-1, 3), np.asarray(faces, np.int64).reshape(-1, 3),
                np.asarray(tri_idx, np.int64).reshape(-1, 3))

    @classmethod
    def from_arrays(cls, vertexes, faces, path="<arrays>"):
        o = cls.__new__(cls)
        o.path = path
        o.vertexes = np.ascontiguousarray(vertexes, np.float64).reshape(-1, 3)
        o.faces = np.ascontiguousarray(faces, np.int64).reshape(-1, 3)
        o._finish()
        return o

    def _finish(self, tri_idx=None, strict=False):
        """tri_idx: resolved vertex index of every triangle corner when it differs from `faces` (the reference stores
        'f 0 1 2' as (-1, 0, 1) and looks the -1 up among the vertices read so far).  strict (files only): a zero-area
        triangle raises ZeroDivisionError like the reference's normalize (vector.py:172-173: v * (1 / v.size()))."""
        idx = self.faces if tri_idx is None else tri_idx
        self.triangles = self.vertexes[idx] if len(idx) else np.zeros((0, 3, 3))
        a = self.triangles[:, 1] - self.triangles[:, 0]
        b = self.triangles[:, 2] - self.triangles[:, 0]
        cx = a[:, 1] * b[:, 2] - a[:, 2] * b[:, 1]               # vector.py:143-146
        cy = a[:, 2] * b[:, 0] - a[:, 0] * b[:, 2]
        cz = a[:, 0] * b[:, 1] - a[:, 1] * b[:, 0]
        with np.errstate(all="ignore"):
            size = np.sqrt(cx * cx + cy * cy + cz * cz)          # vector.py:109-111
            if strict and bool((size == 0.0).any()):
                raise ZeroDivisionError("float division by zero")
            inv = 1.0 / size                                     # vector.py:172-173: v * (1 / size)
            self.normals = np.stack([cx * inv, cy * inv, cz * inv], axis=1)
        self.areas = size / 2                                    # vector.py:164-165


class Scene:
    """`.sdl` scene (scene_reader.py:107-188)."""

    def __init__(self, path=None, quiet=True):
        self.eye = self.width = self.height = self.ortho = self.background = self.ambient = None
        self.light_obj = self.light_color = self.npaths = self.tonemapping = self.seed = None
        self.objects = []
        self.path = path
        if path is not None:
            self.read_scene(path, quiet)

    def read_scene(self, path, quiet=True):
        try:
            lines = _clean_lines(path)
        except OSError as e:
            print("Could not open file!")
            raise e
        for line in lines:
            tok = _tokens(line)
            if not tok:
                continue
            k = tok[0]
            if k == "eye":
                self.eye = [float(t) for t in tok[1:4]]
            elif k == "size":
                self.width, self.height = int(tok[1]), int(tok[2])
            elif k == "ortho":
                self.ortho = [float(t) for t in tok[1:5]]
            elif k == "background":
                self.background = [float(t) for t in tok[1:4]]
            elif k == "ambient":
                self.ambient = float(tok[1])
            elif k == "light":
                self.light_obj = Obj(join(dirname(path), tok[1]), quiet)
                self.light_color = [float(t) for t in tok[2:6]]
            elif k == "npaths":
                self.npaths = int(tok[1])
            elif k == "tonemapping":
                self.tonemapping = float(tok[1])
            elif k == "seed":
                self.seed = int(tok[1])
            elif k == "object":
                d = {"geometry": Obj(join(dirname(path), tok[1]), quiet)}
                for name, t in zip(("red", "green", "blue", "ka", "kd", "ks", "kt", "n"), tok[2:10]):
                    d[name] = float(t)
                if len(d) != 9:
                    raise IndexError("list index out of range")  # reference indexes tokens[2..9]
                self.objects.append(d)
            elif k == "output":
                self.output = join(dirname(path), tok[1])
            else:
                print(f"Scene {path}\n\tSkipping command '{k}'! Parameters: {tok[1:]}")

    def __repr__(self):
        return (f"<Scene eye={self.eye} size={self.width}x{self.height} ortho={self.ortho} ambient={self.ambient} "
                f"objects={len(self.objects)} npaths={self.npaths} seed={self.seed}>")

    # ---- boundary: flatten into the arrays of ptb_scene_desc (include/ptb.h)
    def flatten(self, refresh=False, pin=False):
        """-> dict of the packed arrays of ptb_scene_desc.  The result is cached on the scene (building it copies
        every triangle once: ~1 s for 10 M triangles); pass refresh=True after editing the scene in place.
        pin=True moves the big arrays into page-locked host memory (once; needs a CUDA device), so that every
        Renderer(scene) uploads them at full PCIe / C2C rate instead of through the driver's staging buffer."""
        if refresh or getattr(self, "_flat", None) is None:
            self._flat = self._flatten()
            self._pinned = None
        if pin and getattr(self, "_pinned", None) is None:
            import torch
            self._pinned = {}
            for k in ("verts", "normals", "tri_obj", "light_areas"):
                t = torch.from_numpy(self._flat[k]).pin_memory()
                self._pinned[k] = t                      # owns the page-locked allocation behind the numpy view
                self._flat[k] = t.numpy()
        return self._flat

    def _flatten(self):
        if self.light_obj is None or not self.objects:
            raise ValueError("scene needs a light mesh and at least one object")
        geos = [o["geometry"] for o in self.objects]
        verts = np.concatenate([g.triangles for g in geos] + [self.light_obj.triangles]).astype(np.float64)
        normals = np.concatenate([g.normals for g in geos] + [self.light_obj.normals]).astype(np.float64)
        tri_obj = np.concatenate([np.full(len(g.triangles), i, np.int32) for i, g in enumerate(geos)] +
                                 [np.full(len(self.light_obj.triangles), len(geos), np.int32)])
        mat = np.asarray([[o[k] for k in ("red", "green", "blue", "ka", "kd", "ks", "kt", "n")] for o in self.objects],
                         np.float64)
        return {
            "verts": np.ascontiguousarray(verts), "normals": np.ascontiguousarray(normals),
            "tri_obj": np.ascontiguousarray(tri_obj), "mat": np.ascontiguousarray(mat),
            "light_areas": np.ascontiguousarray(self.light_obj.areas, np.float64),
            "n_tri_obj": int(sum(len(g.triangles) for g in geos)),
            "eye": np.asarray(self.eye, np.float64), "ortho": np.asarray(self.ortho, np.float64),
            "width": self.width, "height": self.height, "ambient": float(self.ambient),
            "light_color": np.asarray(self.light_color[:3], np.float64),
            "npaths": self.npaths, "seed": self.seed, "tonemapping": self.tonemapping, "background": self.background,
        }


CORNELL_SDL = join(dirname(__file__), "scenes", "cornell", "cornellroom.sdl")
