"""Host logic without a GPU: the product's scene loader against the golden dump of the reference's
`Scene`, grammar edge cases, and the C-ABI library's export table."""
import ctypes
import os
import re

import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import _lib
from oracle import scene_port

from conftest import GOLDEN, ROOT


def test_cornell_matches_reference_scene():
    ref = np.load(os.path.join(GOLDEN, "cornell_scene.npz"))
    sc = ptb.Scene(ptb.CORNELL_SDL)
    flat = sc.flatten()
    for k in ("verts", "normals", "tri_obj", "mat", "light_areas", "eye", "ortho", "light_color"):
        assert np.array_equal(ref[k], flat[k]), k      # bit-exact
    for k in ("n_tri_obj", "width", "height", "ambient", "npaths", "seed"):
        assert ref[k] == flat[k], k
    assert sc.tonemapping == 1.0 and sc.background == [0.0, 0.0, 0.0]
    assert sc.output.endswith("cornell.pnm")
    assert [len(o["geometry"].triangles) for o in sc.objects] == [2, 2, 2, 2, 2, 10, 10]


def test_loader_agrees_with_oracle_port():
    a = ptb.Scene(ptb.CORNELL_SDL).flatten()
    b = scene_port.load_scene(ptb.CORNELL_SDL)
    for k in ("verts", "normals", "tri_obj", "mat", "light_areas"):
        assert np.array_equal(a[k], b[k])


def test_obj_grammar(tmp_path, capsys):
    p = tmp_path / "m.obj"
    p.write_text("# comment\n  v 0 0 0\nv 1 0 0 # trailing\nv\t1 1 0\nv 0 1 0\n"
                 "vn 0 0 1\nf 1 2 3 4\nf -4 -3 -2\n\n")
    o = ptb.Obj(str(p))
    assert o.faces.tolist() == [[0, 1, 2], [0, 2, 3], [0, 1, 2]]        # fan + negative indices
    assert np.allclose(o.normals, [[0, 0, 1]] * 3) and np.allclose(o.areas, 0.5)
    assert "Skipping command 'vn'" in capsys.readouterr().out
    # a 4-coordinate vertex next to 3-coordinate ones: the reference's V.__add__ refuses (vector.py:63-72)
    from pathtracerpython_b200.scene import WrongDimensionError
    m = tmp_path / "mixed.obj"
    m.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0 1.0\nf 1 2 3\n")
    with pytest.raises(WrongDimensionError):
        ptb.Obj(str(m))
    q = tmp_path / "slash.obj"
    q.write_text("v 0 0 0\nv 1 0 0\nv 0 1 0\nf 1/1 2/2 3/3\n")
    with pytest.raises(ValueError):
        ptb.Obj(str(q))


def test_sdl_grammar(tmp_path, capsys):
    (tmp_path / "t.obj").write_text("v 0 0 -1\nv 1 0 -1\nv 0 1 -1\nf 1 2 3\n")
    s = tmp_path / "s.sdl"
    s.write_text("eye 0 0 1\nsize 4 4\northo -1 -1 1 1\nambient 0.1\nbogus 1 2\nlight t.obj 1 1 1\n"
                 "light t.obj 0.5 0.5 0.5\nobject t.obj 1 0 0 0.3 0.7 0 0 5\n")
    sc = ptb.Scene(str(s))
    assert sc.light_color == [0.5, 0.5, 0.5]                             # last `light` wins
    assert "Skipping command 'bogus'" in capsys.readouterr().out
    assert sc.npaths is None and sc.seed is None
    with pytest.raises(OSError):
        ptb.Scene(str(tmp_path / "missing.sdl"))
    assert "Could not open file!" in capsys.readouterr().out


def test_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "ptb.h")).read()
    declared = set(re.findall(r"\b(ptb_[a-z_0-9]+)\s*\(", hdr))
    assert declared == set(_lib.EXPORTS), declared ^ set(_lib.EXPORTS)
    L = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(L, name), name
    assert L.ptb_version() >= 100


def test_struct_layouts_match_header():
    # sizes the C side was compiled with (see include/ptb.h)
    assert ctypes.sizeof(_lib.SceneDesc) == 16 + 5 * 8 + 11 * 8 + 8
    assert ctypes.sizeof(_lib.Pixels) == 32
    assert ctypes.sizeof(_lib.RenderParams) == 64
    assert ctypes.sizeof(_lib.PathRecords) == 64
    assert ctypes.sizeof(_lib.Counters) == 80


def test_pixel_sharding_partition():
    L = _lib.lib()
    for (w, h, tile, g) in [(40, 40, 64, 3), (1024, 1024, 4096, 8), (37, 11, 5, 4)]:
        total = 0
        for s in range(g):
            px = _lib.Pixels(w, h, s, g, tile, 0, None)
            total += L.ptb_pixel_count(ctypes.byref(px))
        assert total == w * h


def test_no_cpu_fallback_without_cuda():
    import torch
    if torch.cuda.is_available():
        pytest.skip("CUDA present")
    with pytest.raises(ptb.PtbError):
        ptb.Renderer(ptb.CORNELL_SDL)
