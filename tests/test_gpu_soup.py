"""GPU parity on geometry that is NOT the Cornell box: random triangle soups (intersecting, non-coplanar,
slanted normals, several objects with different kd / ks, a light of three unequal triangles) with the corner
cases the reference's arithmetic has an opinion on - a zero-area triangle (NaN normal: never hit,
utils.py:117-121), a sliver, an exact duplicate (first wins, main.py:113-118), a coplanar quad.  Both traversal
back ends, both precisions, against the C oracle."""
import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200.scene import Obj, Scene
from oracle import oracle

from conftest import record_divergence

pytestmark = pytest.mark.gpu

ATOL, RTOL = 1e-4, 1e-3
LIMIT_SOUP = 4e-4      # divergent path-bounces; measured on B200: 0 - 82 ppm (LBVH), 82 ppm (plane list); profiles/r02a_divergence_measured.json


def soup_scene(n_tri, seed, n_obj=5, spread=2.2, size=0.9):
    rng = np.random.default_rng(seed)
    s = Scene()
    s.eye, s.width, s.height = [0.3, 0.2, 7.0], 64, 64
    s.ortho, s.background, s.ambient = [-2.5, -2.5, 2.5, 2.5], [0.0, 0.0, 0.0], 0.4
    s.light_color, s.npaths, s.tonemapping, s.seed = [1.0, 0.9, 0.8], 1, 1.0, 1
    # light: three triangles of very different areas above the soup, tilted
    lv = np.array([[-1.0, 3.0, -1.0], [1.0, 3.2, -1.0], [1.0, 3.0, 1.0], [-1.0, 2.8, 1.0], [0.0, 3.5, 0.0], [0.2, 3.5, 0.1],
                   [0.1, 3.6, 0.3]])
    s.light_obj = Obj.from_arrays(lv, np.array([[0, 1, 2], [0, 2, 3], [4, 5, 6]]))
    per = n_tri // n_obj
    for k in range(n_obj):
        m = per if k < n_obj - 1 else n_tri - per * (n_obj - 1)
        c = rng.uniform(-spread, spread, (m, 1, 3)) * np.array([1.0, 0.8, 1.0])
        tri = c + rng.normal(0.0, size * rng.uniform(0.05, 1.0, (m, 1, 1)), (m, 3, 3))
        if k == 0:
            tri[0, 2] = tri[0, 1]                                  # zero area: two equal vertices
            tri[1, 2] = tri[1, 0] + (tri[1, 1] - tri[1, 0]) * 0.5 + 1e-7     # sliver
            tri[3] = tri[2]                                        # exact duplicate of the previous triangle
            q = np.array([[-2.0, -2.4, -2.0], [2.0, -2.4, -2.0], [2.0, -2.4, 2.0], [-2.0, -2.4, 2.0]])
            tri[4] = q[[0, 2, 1]]; tri[5] = q[[0, 3, 2]]           # coplanar quad (a floor), normal +y
        verts = tri.reshape(-1, 3)
        faces = np.arange(3 * m).reshape(-1, 3)
        kd, ks = rng.uniform(0.2, 0.8), rng.uniform(0.0, 0.5)
        s.objects.append({"geometry": Obj.from_arrays(verts, faces), "red": rng.uniform(0.2, 1), "green": rng.uniform(0.2, 1),
                          "blue": rng.uniform(0.2, 1), "ka": rng.uniform(0, 0.3), "kd": kd, "ks": ks, "kt": 0.0, "n": 1.0})
    return s


def _np(rec):
    return {k: v.cpu().numpy() for k, v in rec.items()}


def _sig(rec, ref):
    same = (rec["tri"] == ref["tri"]) & (rec["lobe"] == ref["lobe"]) & (rec["occ"] == ref["occ"]) & (rec["leak"] == ref["leak"])
    return np.logical_and.accumulate(same, axis=2)


def check_against_oracle(scene, accel, res, spp, nb, seed, max_div32):
    with np.errstate(invalid="ignore", divide="ignore"):
        flat = scene.flatten()
    orc = oracle.Oracle(flat, res, res)
    want, _, _ = orc.primary()
    assert (want >= 0).mean() > 0.3
    r = ptb.Renderer(scene, device=0, accel=accel)
    for prec in (ptb.PREC_FP32, ptb.PREC_FP64):
        tri, _ = r.trace_primary(res, res, precision=prec)
        assert np.array_equal(tri.cpu().numpy(), want), (accel, prec)
    ref = orc.trace(spp, nb, seed)
    live = ref["tri"] > -2
    assert (ref["tri"][:, :, nb - 1] >= 0).mean() > 0.05            # paths survive to the last bounce
    _, rec = r.trace_paths(spp, nb, seed=seed, width=res, height=res, precision=ptb.PREC_FP64, records=True)
    rec = _np(rec)
    ok = _sig(rec, ref)
    assert (~ok & live).sum() / live.sum() < 1e-4
    assert np.abs(rec["contrib64"] - ref["contrib"])[ok].max() < 1e-9
    acc, rec = r.trace_paths(spp, nb, seed=seed, width=res, height=res, records=True)
    rec = _np(rec)
    ok = _sig(rec, ref)
    div = record_divergence(f"soup {flat['verts'].shape[0]} triangles, accel {accel}, {res}x{res}x{spp}spp b{nb}", (~ok & live).sum(), live.sum(), max_div32)
    assert div < max_div32, div
    assert (np.abs(rec["contrib"] - ref["contrib"]) <= ATOL + RTOL * np.abs(ref["contrib"]))[ok].all()
    a = acc.cpu().numpy()
    assert np.isfinite(a).all()
    r.close()
    return div


@pytest.mark.parametrize("accel", [ptb.ACCEL_BRUTE, ptb.ACCEL_LBVH])
def test_soup_200_both_back_ends_vs_oracle(accel):
    check_against_oracle(soup_scene(200, 11), accel, 96, 3, 4, 77, LIMIT_SOUP)


def test_soup_512_planes_fill_shared_memory():
    """512 triangles, no two coplanar: 512 plane records = 49 168 B of dynamic shared memory, more than a kernel
    gets without opting in."""
    sc = soup_scene(509, 14)
    geo = sc.objects[0]["geometry"]
    v = geo.vertexes.copy()
    v[0:18] += np.random.default_rng(1).normal(0.0, 0.2, (18, 3))          # undo the degenerate / duplicate / coplanar cases
    sc.objects[0]["geometry"] = Obj.from_arrays(v, geo.faces)
    lv = sc.light_obj.vertexes.copy()
    lv[3, 1] = 2.7                                                         # the light's quad is no longer planar
    sc.light_obj = Obj.from_arrays(lv, sc.light_obj.faces)
    flat = sc.flatten(refresh=True)
    assert flat["verts"].shape[0] == 512
    r = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_BRUTE)
    want, _, _ = oracle.Oracle(flat, 64, 64).primary()
    tri, _ = r.trace_primary(64, 64)
    assert np.array_equal(tri.cpu().numpy(), want)
    ref = oracle.Oracle(flat, 32, 32).trace(2, 3, 2)
    _, rec = r.trace_paths(2, 3, seed=2, width=32, height=32, records=True)
    rec = _np(rec)
    ok = _sig(rec, ref)
    live = ref["tri"] > -2
    assert record_divergence("soup 512 planes in shared memory, 32x32x2spp b3", (~ok & live).sum(), live.sum(), 6e-4) < 6e-4   # 1 of 5897 measured
    assert (np.abs(rec["contrib"] - ref["contrib"]) <= ATOL + RTOL * np.abs(ref["contrib"]))[ok].all()
    r.close()


def test_soup_5000_lbvh_vs_oracle():
    check_against_oracle(soup_scene(5000, 12, n_obj=7, size=0.35), ptb.ACCEL_LBVH, 96, 2, 3, 5, LIMIT_SOUP)


def test_soup_back_ends_agree_with_each_other():
    sc = soup_scene(300, 13)
    rb = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_BRUTE)
    rl = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_LBVH)
    tb, _ = rb.trace_primary(200, 200)
    tl, _ = rl.trace_primary(200, 200)
    assert np.array_equal(tb.cpu().numpy(), tl.cpu().numpy())
    a, _ = rb.trace_paths(8, 4, seed=3, width=80, height=80)
    b, _ = rl.trace_paths(8, 4, seed=3, width=80, height=80)
    a, b = a.cpu().numpy(), b.cpu().numpy()
    close = np.abs(a - b) <= 1e-3 + 1e-3 * np.abs(a)
    assert close.mean() > 0.99
    rb.close(); rl.close()
