"""INTEGRATION.md section B is the ctypes stub a maintainer of the reference would paste into main.py.  This runs that
block exactly as printed - against the reference's OWN scene_reader.Scene object when oracle/_ref holds the copy of the
reference (oracle/make_ref.py; it travels to the GPU box), else against the product's Scene wrapped to look like it -
and holds its image equal to the shipped host package's."""
import os
import re
import sys

import numpy as np
import pytest

import pathtracerpython_b200 as ptb

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _stub_source():
    text = open(os.path.join(ROOT, "INTEGRATION.md")).read()
    m = re.search(r"# --- begin stub.*?\n(.*?)# --- end stub ---", text, re.S)
    assert m, "INTEGRATION.md: stub markers not found"
    return m.group(1)


class _V:                        # what flatten() touches of vector.V
    def __init__(self, c):
        self.coords = tuple(float(x) for x in c)


class _Geo:
    def __init__(self, obj):
        self.triangles = [tuple(_V(v) for v in t) for t in obj.triangles]
        self.normals = [_V(n) for n in obj.normals]
        self.areas = [float(a) for a in obj.areas]


def _reference_like(scene):
    class S:
        pass
    s = S()
    for k in ("eye", "width", "height", "ortho", "ambient", "light_color"):
        setattr(s, k, getattr(scene, k))
    s.light_obj = _Geo(scene.light_obj)
    s.objects = [dict(o, geometry=_Geo(o["geometry"])) for o in scene.objects]
    return s


def test_integration_stub_runs_and_matches_the_host_package(monkeypatch):
    ref_dir = os.path.join(ROOT, "oracle", "_ref", "pathtracerpython")
    if os.path.exists(os.path.join(ref_dir, "scene_reader.py")):
        monkeypatch.syspath_prepend(ref_dir)
        import scene_reader                                     # the reference's own parser, unmodified
        import contextlib
        import io
        with contextlib.redirect_stdout(io.StringIO()):
            scene = scene_reader.Scene(os.path.join(ref_dir, "objs", "cornellroom.sdl"))
        sys.modules.pop("scene_reader", None)
    else:
        scene = _reference_like(ptb.Scene(ptb.CORNELL_SDL))
    ns = {}
    monkeypatch.chdir(ROOT)                                     # the stub loads "pathtracerpython_b200/libptb.so"
    exec(compile(_stub_source(), "INTEGRATION.md", "exec"), ns)
    im = ns["render_with_libptb"](scene, 4, 3, seed=9)
    want = ptb.render(ptb.CORNELL_SDL, 4, 3, seed=9)
    assert im.size == (40, 40)
    assert np.array_equal(np.asarray(im), np.asarray(want))
