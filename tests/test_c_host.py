"""The C ABI used from a host program in plain C (tests/c_host/ptb_host.c): no Python or torch on the path.  The CPU
test compiles and links it against libptb.so (every symbol it uses resolves); the GPU test runs it on the Cornell box
and on a tessellated one (LBVH) and compares its output, byte for byte, with the Python host's."""
import os
import struct
import subprocess

import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import _lib

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SRC = os.path.join(ROOT, "tests", "c_host", "ptb_host.c")
CUDA = os.environ.get("CUDA_HOME", "/usr/local/cuda")


def build(tmp_path):
    exe = str(tmp_path / "ptb_host")
    libdir = os.path.dirname(_lib.LIB_PATH)
    cmd = ["gcc", "-std=c99", "-O1", "-Wall", "-o", exe, SRC, "-I", os.path.join(ROOT, "include"), "-I", os.path.join(CUDA, "include"),
           "-L", libdir, "-l:libptb.so", "-L", os.path.join(CUDA, "lib64"), "-lcudart", f"-Wl,-rpath,{libdir}",
           f"-Wl,-rpath,{os.path.join(CUDA, 'lib64')}"]
    out = subprocess.run(cmd, capture_output=True, text=True)
    assert out.returncode == 0, out.stderr
    return exe


def write_scene(path, flat, accel):
    v, nrm, tobj, mat, areas = (np.ascontiguousarray(flat[k], dt) for k, dt in
                                (("verts", np.float64), ("normals", np.float64), ("tri_obj", np.int32), ("mat", np.float64),
                                 ("light_areas", np.float64)))
    with open(path, "wb") as f:
        f.write(struct.pack("<6i", v.shape[0], int(flat["n_tri_obj"]), mat.shape[0], areas.shape[0], accel, 0))
        for a in (v, nrm, mat, areas):
            f.write(a.tobytes())
        tail = list(flat["eye"]) + list(flat["ortho"]) + [flat["ambient"]] + list(flat["light_color"][:3])
        f.write(np.asarray(tail, np.float64).tobytes())
        f.write(tobj.tobytes())


def test_c_host_compiles_and_links(tmp_path):
    exe = build(tmp_path)
    assert os.path.exists(exe)
    out = subprocess.run([exe], capture_output=True, text=True)            # no arguments: usage, before any CUDA call
    assert out.returncode == 1 and "usage" in out.stderr


@pytest.mark.gpu
@pytest.mark.parametrize("tess", [0, 6])
def test_c_host_matches_python_host(tmp_path, tess):
    from pathtracerpython_b200 import synth
    exe = build(tmp_path)
    scene = ptb.Scene(ptb.CORNELL_SDL)
    if tess:
        scene = synth.tessellate(scene, tess)
    flat = scene.flatten()
    W, H, spp, nb, seed = 48, 40, 3, 3, 11
    write_scene(str(tmp_path / "scene.bin"), flat, ptb.ACCEL_AUTO)
    out = subprocess.run([exe, str(tmp_path / "scene.bin"), str(tmp_path / "out.bin"), str(W), str(H), str(spp), str(nb), str(seed)],
                         capture_output=True, text=True, timeout=300)
    assert out.returncode == 0, out.stdout + out.stderr
    assert "ptb_host ok" in out.stdout
    raw = open(str(tmp_path / "out.bin"), "rb").read()
    n = W * H
    tri = np.frombuffer(raw, np.int32, n)
    accum = np.frombuffer(raw, np.float32, 4 * n, 4 * n).reshape(n, 4)
    rgb = np.frombuffer(raw, np.uint8, 3 * n, 20 * n).reshape(H, W, 3)
    r = ptb.Renderer(scene, device=0)
    t2, _ = r.trace_primary(W, H)
    a2, _ = r.trace_paths(spp, nb, seed=seed, width=W, height=H, samples_per_batch=spp)
    assert np.array_equal(tri, t2.cpu().numpy())
    assert np.array_equal(accum, a2.cpu().numpy())
    assert np.array_equal(rgb, r.image(a2, W, H).cpu().numpy())
    r.close()
