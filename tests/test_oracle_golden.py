"""Pins the oracle (oracle/pt_oracle.c, oracle/scene_port.py, oracle/philox.py) against the golden
vectors produced by running the unmodified reference (tests/golden/make_golden.py)."""
import os

import numpy as np
import pytest

from oracle import oracle, philox, scene_port

from conftest import GOLDEN

SDL = os.path.join(os.path.dirname(GOLDEN), "..", "pathtracerpython_b200", "scenes", "cornell", "cornellroom.sdl")


def g(name):
    return np.load(os.path.join(GOLDEN, name))


@pytest.fixture(scope="module")
def flat():
    return scene_port.load_scene(os.path.abspath(SDL))


@pytest.fixture(scope="module")
def orc(flat):
    return oracle.Oracle(flat)


def test_philox_kat():
    # Random123 known answers
    kats = [((0, 0, 0, 0), (0, 0), (0x6627e8d5, 0xe169c58d, 0xbc57ac4c, 0x9b00dbd8)),
            ((0xffffffff,) * 4, (0xffffffff,) * 2, (0x408f276d, 0x41c83b0e, 0xa20bc7c6, 0x6d5451fd)),
            ((0x243f6a88, 0x85a308d3, 0x13198a2e, 0x03707344), (0xa4093822, 0x299f31d0),
             (0xd16cfe09, 0x94fdcceb, 0x5001e420, 0x24126ea1))]
    for ctr, key, want in kats:
        assert philox.philox4x32_10(ctr, key) == want
        assert tuple(int(x) for x in oracle.philox(ctr, key)) == want


def test_scene_port_matches_reference_scene(flat):
    ref = g("cornell_scene.npz")
    for k in ("verts", "normals", "tri_obj", "mat", "light_areas", "eye", "ortho", "light_color"):
        assert np.array_equal(ref[k], flat[k]), k          # bit-exact: same Python-float operations
    for k in ("n_tri_obj", "width", "height", "ambient", "npaths", "seed"):
        assert ref[k] == flat[k], k


def test_intersect_known_answers(flat):
    kat = g("kat_intersect.npz")
    n_hit = 0
    for ray, ti, hit, pt in zip(kat["rays"], kat["tri"], kat["hit"], kat["point"]):
        p = oracle.intersect_one(ray[:3], ray[3:], flat["verts"][ti])
        assert (p is not None) == bool(hit)
        if hit:
            n_hit += 1
            assert np.allclose(p, pt, rtol=0, atol=1e-11)
    assert n_hit > 100


def test_primary_ids_bit_exact(flat):
    gold = g("cornell_primary.npz")
    tri, _, _ = oracle.Oracle(flat).primary()
    assert np.array_equal(tri, gold["tri_40"])
    for res in (512, 1024):
        o = oracle.Oracle(flat, res, res)
        tri, _, pt = o.primary(gold[f"pixels_{res}"])
        assert np.array_equal(tri, gold[f"tri_{res}"])
        hit = tri >= 0
        assert np.abs(pt[hit] - gold[f"point_{res}"][hit]).max() < 1e-11


def test_paths_match_reference(orc):
    gold = g("cornell_paths_40.npz")
    out = orc.trace(int(gold["spp"]), int(gold["bounces"]), int(gold["seed"]))
    for k in ("tri", "lobe", "occ", "leak"):
        assert np.array_equal(out[k], gold[k]), k          # every discrete outcome identical
    for k in ("point", "contrib", "k_after", "radiance"):
        assert np.abs(out[k] - gold[k]).max() < 1e-12, k
    assert np.array_equal(orc.make_image(out["radiance"]), gold["image"])


def test_misc_known_answers(flat):
    kat = g("kat_misc.npz")
    o = oracle.Oracle(flat, 24, 24)
    assert np.array_equal(o.make_image(kat["img_radiance"]), kat["img_u8"])
    # rotate (main.py:148-162), pick_random_triangle + sample_random_pt (utils.py:21-46), make_rays (utils.py:55-69):
    # the reference's own outputs for these four functions
    assert np.abs(oracle.rotate_y(kat["rot_angle"], kat["rot_in"]) - kat["rot_out"]).max() < 1e-14
    pos = np.array([[pix, 3, 1, 0] for pix in range(len(kat["light_pick"]))], np.int32)
    pick, pt = oracle.Oracle(flat).light_sample(9, pos)
    assert np.array_equal(pick, kat["light_pick"])
    assert np.abs(pt - kat["light_pt"]).max() < 1e-14
    assert len(set(kat["light_pick"].tolist())) == 2          # both light triangles are drawn
    dirs = oracle.Oracle(flat, 7, 7).primary_dirs(np.arange(49))
    assert np.array_equal(dirs, kat["ray_dirs"])              # FP64 linspace, bit for bit


def test_trace_does_not_depend_on_how_threads_share_the_work(flat):
    """pto_trace splits few pixels per (pixel, sample) and many per pixel: same records, same radiance."""
    o = oracle.Oracle(flat, 64, 64)
    px = np.arange(100, 110)
    few = o.trace(3, 3, 5, pixels=px)
    n = oracle.num_threads()
    many = o.trace(3, 3, 5, pixels=np.concatenate([px, np.arange(200, 200 + 64 * n)]))
    for k in ("tri", "lobe", "occ", "leak", "contrib", "k_after", "radiance"):
        assert np.array_equal(few[k], many[k][:len(px)]), k
    assert oracle.set_threads(0) == oracle.host_cores()
