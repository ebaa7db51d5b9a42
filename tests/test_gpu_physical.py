"""GPU tests of physical mode (PTB_FLAG_PHYSICAL).  The reference has no such mode, so
it is validated on its own terms: an analytic direct-lighting integral, energy bounds, Russian-roulette
unbiasedness, a transmissive pane, and agreement between the two traversal back ends."""
import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200.scene import Obj, Scene

pytestmark = pytest.mark.gpu

PHYS = ptb.FLAG_PHYSICAL


def quad(p0, p1, p2, p3):
    """two triangles (p0 p1 p2), (p0 p2 p3); normal = (p1-p0) x (p2-p0)"""
    return Obj.from_arrays(np.array([p0, p1, p2, p3], np.float64), np.array([[0, 1, 2], [0, 2, 3]]))


def make_scene(objects, light, light_color=(1.0, 1.0, 1.0), size=32, window=1.0):
    s = Scene()
    s.eye, s.width, s.height = [0.0, 0.0, 6.0], size, size
    s.ortho, s.background, s.ambient = [-window, -window, window, window], [0.0, 0.0, 0.0], 0.0
    s.light_obj, s.light_color = light, list(light_color)
    s.npaths, s.tonemapping, s.seed = 1, 1.0, 1
    for geo, rgb, kd, ks, kt, n in objects:
        s.objects.append({"geometry": geo, "red": rgb[0], "green": rgb[1], "blue": rgb[2], "ka": 0.0, "kd": kd, "ks": ks,
                          "kt": kt, "n": n})
    return s


def radiance(r, spp, bounces, flags=PHYS, seed=3, **kw):
    acc, _ = r.trace_paths(spp, bounces, seed=seed, flags=flags, **kw)
    a = acc.cpu().numpy().astype(np.float64)
    return a[:, :3] / a[:, 3:4]


def test_direct_lighting_matches_the_analytic_integral():
    """Diffuse wall z = -4 facing the camera, square emitter in the plane z = -2 facing the wall, one bounce:
    L_o(P) = albedo * kd / pi * integral_A Le cos_s cos_l / r^2 dA, evaluated here by quadrature."""
    zw, zl, h, albedo, kd, Le = -4.0, -2.0, 0.4, (0.8, 0.6, 0.4), 0.9, 2.0
    wall = quad([-3, -3, zw], [3, -3, zw], [3, 3, zw], [-3, 3, zw])             # normal +z
    light = quad([-h, -h, zl], [-h, h, zl], [h, h, zl], [h, -h, zl])            # normal -z (two-sided anyway)
    sc = make_scene([(wall, albedo, kd, 0.0, 0.0, 1.0)], light, (Le, Le, Le), size=24)
    r = ptb.Renderer(sc, device=0)
    rad = radiance(r, 2048, 1)
    tri, t = r.trace_primary()
    tri = tri.cpu().numpy()
    W = H = 24
    xs = np.linspace(-1, 1, W); ys = np.linspace(-1, 1, H)
    eye = np.array(sc.eye)
    g = (np.arange(400) + 0.5) / 400 * 2 * h - h
    gx, gy = np.meshgrid(g, g, indexing="ij")
    dA = (2 * h / 400) ** 2
    checked, rel = 0, []
    for ix in range(W):
        for iy in range(H):
            idx = ix * H + iy
            if tri[idx] < 0 or tri[idx] >= 2:          # the emitter (ids 2, 3) hides the wall in the centre
                continue
            d = np.array([xs[ix], ys[iy], 0.0]) - eye
            P = eye + d * ((zw - eye[2]) / d[2])
            vx, vy, vz = gx - P[0], gy - P[1], zl - P[2]
            r2 = vx * vx + vy * vy + vz * vz
            E = (Le * (vz / np.sqrt(r2)) ** 2 / r2).sum() * dA          # cos_s = cos_l = vz / r
            want = np.array(albedo) * kd / np.pi * E
            rel.append((rad[idx] - want) / want)
            checked += 1
    rel = np.array(rel)
    assert checked > 300
    assert np.abs(rel.mean(axis=0)).max() < 0.01, rel.mean(axis=0)      # unbiased: mean error over the frame < 1 %
    assert np.abs(rel).max() < 0.12                                    # per pixel: 2048 x 3 light samples
    # pixels that see the emitter directly get exactly Le
    lit = (tri >= 2)
    assert lit.any() and np.allclose(rad[lit], Le)
    r.close()


def test_closed_room_energy_bounds_and_roulette_unbiased():
    r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    a = radiance(r, 512, 6, width=48, height=48)
    b = radiance(r, 512, 6, flags=PHYS | ptb.FLAG_NO_ROULETTE, width=48, height=48)
    assert np.isfinite(a).all() and (a >= 0).all() and (b >= 0).all()
    # albedo <= 1 everywhere and the only emitter has radiance 1: no pixel's expectation can exceed it
    assert a.mean() < 1.0 and a.max() < 3.0
    assert abs(a.mean() - b.mean()) < 0.02 * b.mean(), (a.mean(), b.mean())
    # colour bleeding exists in physical mode (RGB throughput): the red wall tints its neighbourhood
    assert a[:, 0].mean() > a[:, 2].mean()
    # deeper paths only add light
    c = radiance(r, 512, 2, width=48, height=48)
    assert c.mean() < a.mean()
    r.close()


def test_transmissive_pane_passes_camera_rays_to_the_emitter():
    pane = quad([-3, -3, -1], [3, -3, -1], [3, 3, -1], [-3, 3, -1])              # normal +z, faces the camera
    light = quad([-3, -3, -3], [-3, 3, -3], [3, 3, -3], [3, -3, -3])
    sc = make_scene([(pane, (1, 1, 1), 0.0, 0.0, 1.0, 1.0)], light, (1.5, 1.5, 1.5), size=16, window=0.5)
    r = ptb.Renderer(sc, device=0)
    one = radiance(r, 4, 1)
    two = radiance(r, 4, 2)
    assert np.allclose(one, 0.0)              # the path ends on the pane: no NEE on a purely transmissive surface
    assert np.allclose(two, 1.5)              # refracted ray (delta lobe) reaches the emitter: emission counted, throughput 1
    # an opaque diffuse pane instead: emission behind it is never seen, only NEE from the far side (none: light is behind)
    sc2 = make_scene([(pane, (1, 1, 1), 1.0, 0.0, 0.0, 1.0)], light, (1.5, 1.5, 1.5), size=16, window=0.5)
    r2 = ptb.Renderer(sc2, device=0)
    assert np.allclose(radiance(r2, 4, 3), 0.0)
    r.close(); r2.close()


def test_lbvh_and_plane_kernels_agree_in_physical_mode():
    sc = ptb.Scene(ptb.CORNELL_SDL)
    rb = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_BRUTE)
    rl = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_LBVH)
    a = radiance(rb, 16, 4, width=64, height=64)
    b = radiance(rl, 16, 4, width=64, height=64)
    close = np.abs(a - b) <= 1e-4 + 1e-3 * np.abs(a)
    assert close.mean() > 0.995                # same stream, same decisions; FP32 edge cases may flip a few paths
    assert abs(a.mean() - b.mean()) < 0.005 * a.mean()
    rb.close(); rl.close()


def test_physical_mode_rejects_reference_only_options():
    r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    with pytest.raises(ptb.PtbError):
        r.trace_paths(1, 1, flags=PHYS, precision=ptb.PREC_FP64)
    with pytest.raises(ptb.PtbError):
        r.trace_paths(1, 1, flags=PHYS, records=True)
    r.close()
