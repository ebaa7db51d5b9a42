"""ptb_obj_parse (libptb.so host code) against the Python restatement of the reference's
.obj grammar (scene_reader.py:29-104): same arrays bit for bit, same skipped-command reports, same errors.
Runs without a GPU: the parser is host code."""
import os
import time

import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import _lib, synth
from pathtracerpython_b200.scene import Obj

HERE = os.path.dirname(os.path.abspath(__file__))
SCENES = os.path.join(os.path.dirname(HERE), "pathtracerpython_b200", "scenes")

pytestmark = pytest.mark.skipif(not os.path.exists(_lib.LIB_PATH), reason="libptb.so not built")


def both(path):
    a = Obj(path, native=False)
    b = Obj(path, native=True)
    return a, b


def same(a, b):
    assert a.vertexes.dtype == b.vertexes.dtype and a.faces.dtype == b.faces.dtype
    assert a.vertexes.shape == b.vertexes.shape and a.faces.shape == b.faces.shape
    assert a.vertexes.tobytes() == b.vertexes.tobytes()
    assert np.array_equal(a.faces, b.faces)
    assert a.triangles.tobytes() == b.triangles.tobytes() and a.normals.tobytes() == b.normals.tobytes()


def test_shipped_meshes_parse_identically():
    n = 0
    for root, _, files in os.walk(SCENES):
        for f in files:
            if f.endswith(".obj"):
                same(*both(os.path.join(root, f)))
                n += 1
    assert n >= 5


from obj_corners import CASES  # noqa: E402

GOLDEN_CORNERS = os.path.join(HERE, "golden", "obj_corners.json")


@pytest.mark.filterwarnings("ignore:invalid value", "ignore:overflow")          # the inf vertices, in Obj._finish
@pytest.mark.parametrize("name", sorted(CASES))
def test_grammar_corners_like_the_reference_parser(tmp_path, capsys, name):
    """Every corner of the grammar against what the REFERENCE's parser did with the same text
    (tests/golden/obj_corners.json, written by tests/golden/make_golden.py from the unmodified scene_reader.Obj): the
    same exception class, or the same vertices (first three coordinates), faces as stored, triangles, normals, areas
    and number of skipped-command reports - from the line-by-line parser and from ptb_obj_parse."""
    import json
    gold = json.load(open(GOLDEN_CORNERS))[name]
    p = tmp_path / "c.obj"
    with open(p, "w", newline="") as f:
        f.write(CASES[name])
    for native in (False, True):
        capsys.readouterr()
        if gold["raises"]:
            with pytest.raises(Exception) as ei:
                Obj(str(p), native=native)
            assert type(ei.value).__name__ == gold["raises"], (native, ei.value)
            continue
        o = Obj(str(p), native=native)
        assert capsys.readouterr().out.count("Skipping command") == gold["skipped"], native
        want_v = np.array([(v + [0.0, 0.0])[:3] for v in gold["vertexes"]], np.float64).reshape(-1, 3)
        assert o.vertexes.tobytes() == want_v.tobytes(), native
        assert o.faces.tolist() == gold["faces"], native
        for k in ("triangles", "normals", "areas"):
            want = np.array(gold[k], np.float64).reshape(getattr(o, k).shape)
            assert np.array_equal(getattr(o, k), want, equal_nan=True), (native, k)      # bit for bit (NaN where the reference has NaN)


def test_corner_goldens_cover_every_case():
    import json
    gold = json.load(open(GOLDEN_CORNERS))
    assert set(gold) == set(CASES)
    assert {g["raises"] for g in gold.values()} >= {None, "WrongDimensionError", "IndexError", "ValueError", "ZeroDivisionError"}


def test_crlf_and_missing_final_newline(tmp_path):
    p = tmp_path / "w.obj"
    p.write_bytes(b"v 0 0 0\r\nv 1 0 0\r\nv 0 1 0\r\nf 1 2 3")
    same(*both(str(p)))


@pytest.mark.parametrize("text, exc", [
    ("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1/1/1 2/2/2 3/3/3\n", ValueError),      # scene_reader.py:69
    ("v 0 0 zero\n", ValueError),
    ("v 0 0 0x10\n", ValueError),
    ("v 0 0 0\nf 1 2 3.0\n", ValueError),
    ("v 0 0 0\n   ", IndexError),                                          # scene_reader.py:31
])
def test_errors_match_the_reference(tmp_path, text, exc):
    p = tmp_path / "e.obj"
    p.write_text(text)
    with pytest.raises(exc):
        Obj(str(p), native=False)
    with pytest.raises(exc):
        Obj(str(p), native=True)


def test_missing_file_is_an_oserror():
    with pytest.raises(OSError):
        Obj("/nonexistent/mesh.obj", native=False)
    with pytest.raises(OSError):
        Obj("/nonexistent/mesh.obj", native=True)


def test_random_decimals_round_like_python(tmp_path):
    rng = np.random.default_rng(5)
    vals = np.concatenate([rng.standard_normal(3000) * 10.0 ** rng.integers(-12, 12, 3000),
                           rng.integers(-10 ** 6, 10 ** 6, 3000).astype(np.float64)])
    fmts = ["%.17g", "%.6f", "%.3e", "%r", "%.15g", "%.20f"]
    lines = []
    for i in range(0, len(vals) - 2, 3):
        f = fmts[(i // 3) % len(fmts)]
        lines.append("v " + " ".join((repr(float(x)) if f == "%r" else f % x) for x in vals[i:i + 3]))
    p = tmp_path / "r.obj"
    p.write_text("\n".join(lines) + "\nf 1 2 3\n")
    same(*both(str(p)))


def test_multithreaded_slices_keep_relative_indices(tmp_path):
    """> 1 MiB so the file is cut into per-thread slices: interleaved vertex / face blocks with negative
    (relative) indices must resolve against the vertices read SO FAR across slice borders."""
    rng = np.random.default_rng(9)
    lines = []
    for k in range(30000):
        for _ in range(3):
            lines.append("v %.9f %.9f %.9f" % tuple(rng.uniform(-5, 5, 3)))
        lines.append("f -3 -2 -1" if k % 2 else "f %d %d %d" % (3 * k + 1, 3 * k + 2, 3 * k + 3))
        if k % 1000 == 0:
            lines.append("o part%d" % k)
    p = tmp_path / "big.obj"
    p.write_text("\n".join(lines) + "\n")
    assert os.path.getsize(p) > (1 << 20)
    a, b = both(str(p))
    same(a, b)
    v1, f1, s1 = _lib.obj_parse(str(p), threads=1)
    v7, f7, s7 = _lib.obj_parse(str(p), threads=7)
    assert v1.tobytes() == v7.tobytes() and np.array_equal(f1, f7) and s1 == s7 and len(s1) == 30
    assert np.array_equal(f1, np.arange(90000).reshape(-1, 3))


def test_writer_round_trips_every_double(tmp_path):
    rng = np.random.default_rng(3)
    v = np.concatenate([rng.standard_normal(3000) * 10.0 ** rng.integers(-30, 30, 3000),
                        np.array([0.0, -0.0, 5e-324, 1.7976931348623157e308, 1e22, 1e23, 0.1, 1 / 3, -1e-5])]).reshape(-1, 3)
    f = rng.integers(0, len(v), (500, 3))
    # a zero-area face raises ZeroDivisionError when read back, like the reference (vector.py:172-173): keep the others
    t = v[f]
    with np.errstate(all="ignore"):
        size = np.sqrt((np.cross(t[:, 1] - t[:, 0], t[:, 2] - t[:, 0]) ** 2).sum(1))
    f = f[size != 0.0]
    assert len(f) > 400
    p = str(tmp_path / "w.obj")
    _lib.obj_write(p, v, f, "round trip")
    for native in (False, True):
        o = Obj(p, native=native)
        assert o.vertexes.tobytes() == v.tobytes() and np.array_equal(o.faces, f)
    assert open(p).readline() == "# round trip\n"
    with pytest.raises(OSError):
        _lib.obj_write("/nonexistent/dir/x.obj", v, f)


def test_tessellated_scene_round_trip_is_fast(tmp_path):
    """write_scene -> Scene(path) of a 200K-triangle config-4 style scene: the native parser gives the arrays
    the generator wrote, and is much faster than the interpreter loop."""
    sc = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 112)                 # 32 x 112^2 = 401 408 triangles
    path = synth.write_scene(sc, str(tmp_path), "tess112.sdl")
    t0 = time.perf_counter()
    s2 = ptb.Scene(path)
    t_native = time.perf_counter() - t0
    f1, f2 = sc.flatten(), s2.flatten()
    assert f1["verts"].shape == f2["verts"].shape and f1["verts"].shape[0] > 400000
    assert np.array_equal(f1["verts"], f2["verts"]) and np.array_equal(f1["tri_obj"], f2["tri_obj"])
    big = max((os.path.join(str(tmp_path), f) for f in os.listdir(str(tmp_path)) if f.endswith(".obj")), key=os.path.getsize)
    t0 = time.perf_counter()
    a = Obj(big, native=False)
    t_py = time.perf_counter() - t0
    t0 = time.perf_counter()
    b = Obj(big, native=True)
    t_c = time.perf_counter() - t0
    same(a, b)
    assert t_c < t_py, (t_c, t_py)
    print(f"scene load {t_native:.2f} s; largest mesh {len(a.faces)} triangles: python {t_py:.2f} s, native {t_c:.2f} s")
