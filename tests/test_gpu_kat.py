"""Known answers of the device functions that have no seam of their own (tier 1) - rotate, the light
triangle pick, the light point, the primary-ray directions - against the REFERENCE's own outputs
(tests/golden/kat_misc.npz, written by tests/golden/make_golden.py from the unmodified main.py / utils.py), through the
C ABI's ptb_kat.  FP64 flavour to 1e-12 (exact for the discrete pick and the ray directions), FP32 flavour - the
functions the product kernels call - within FP32 rounding."""
import os

import numpy as np
import pytest

import pathtracerpython_b200 as ptb

from conftest import GOLDEN

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def rnd():
    r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    yield r
    r.close()


@pytest.fixture(scope="module")
def kat():
    return np.load(os.path.join(GOLDEN, "kat_misc.npz"))


def test_rotate_known_answers(rnd, kat):
    data = np.concatenate([kat["rot_angle"][:, None], kat["rot_in"]], 1)
    out64 = rnd.kat("rotate", data, ptb.PREC_FP64)
    assert np.abs(out64 - kat["rot_out"]).max() < 1e-12
    out32 = rnd.kat("rotate", data, ptb.PREC_FP32)
    assert np.abs(out32 - kat["rot_out"]).max() < 5e-6 * max(1.0, np.abs(kat["rot_out"]).max())


def test_light_sampling_known_answers(rnd, kat):
    n = len(kat["light_pick"])
    pos = np.array([[pix, 3, 1, 0] for pix in range(n)], np.float64)       # make_golden.py: st.at(pix, 3, 1, 0)
    for prec, tol in ((ptb.PREC_FP64, 1e-12), (ptb.PREC_FP32, 2e-6)):
        out = rnd.kat("light", pos, prec, seed=9)
        assert np.array_equal(out[:, 0].astype(np.int32), kat["light_pick"]), prec     # the discrete pick is exact in both
        assert np.abs(out[:, 1:] - kat["light_pt"]).max() < tol * max(1.0, np.abs(kat["light_pt"]).max()), prec


def test_primary_ray_directions_known_answers(rnd, kat):
    out = rnd.kat("ray_dir", np.arange(49, dtype=np.float64), ptb.PREC_FP64, width=7, height=7)
    assert np.array_equal(out, kat["ray_dirs"])                             # FP64 linspace, bit for bit
