"""Command line with the reference's flags (main.py:125-139):

    python -m pathtracerpython_b200 SCENE [--out F] [-r N] [-b N]

`--show-img` / `--show-scene` / `--show-normals` / `--show-screen` / `--show-inter` are accepted and
ignored: they open windows (PIL viewer, OpenGL debug viewer) and there is no display or graphics stack
on this box.  Extras (off by default, the reference ignores these scene keywords, A-Q19):
`--use-npaths` takes -r from the scene's `npaths`, `--seed` overrides the scene's `seed` as Philox key,
`--size W H` overrides `size`, `--physical` selects physical mode (ray queries, true tangent frame, RGB
throughput, refraction for `kt`, Russian roulette; include/ptb.h PTB_FLAG_PHYSICAL), `--use-output` saves to the
scene's `output` file when --out is not given (`.pnm` is written as binary PPM), `--scene-tonemap` applies the scene's
`tonemapping` / `background` keywords instead of make_image's min/max normalisation, `--aov F.npz` dumps the
primary-hit buffers (triangle id, ray parameter, hit point, normal, object id) that the reference's debug
viewer draws (plot.py:28-42, 98-105) - as arrays, there is no window."""
import argparse

import numpy as np

from . import FLAG_PHYSICAL, Renderer, render
from .scene import Scene


def main(argv=None):
    ap = argparse.ArgumentParser(prog="pathtracerpython_b200")
    ap.add_argument("scene", help="SDL scene")
    ap.add_argument("--out", help="Output image")
    ap.add_argument("-r", dest="n_rays", type=int, default=1, help="Number of rays per pixel")
    ap.add_argument("-b", dest="n_bounces", type=int, default=1, help="Number of bounces")
    for flag in ("--show-img", "--show-scene", "--show-normals", "--show-screen", "--show-inter"):
        ap.add_argument(flag, default=False, action="store_true", help="accepted, ignored (no display)")
    ap.add_argument("--use-npaths", action="store_true")
    ap.add_argument("--seed", type=int, default=None)
    ap.add_argument("--size", type=int, nargs=2, default=None)
    ap.add_argument("--physical", action="store_true")
    ap.add_argument("--use-output", action="store_true", help="without --out, save to the scene's `output` file")
    ap.add_argument("--scene-tonemap", action="store_true", help="use the scene's tonemapping/background keywords")
    ap.add_argument("--aov", default=None, help="write primary-hit AOVs to this .npz")
    a = ap.parse_args(argv)
    scene = Scene(a.scene, quiet=False)
    print(f"Number of objects: {len(scene.objects)}")
    print(f"Number of triangles: {sum(len(o['geometry'].triangles) for o in scene.objects)}")
    n_rays = scene.npaths if (a.use_npaths and scene.npaths) else a.n_rays
    w, h = a.size if a.size else (None, None)
    if a.aov:
        r = Renderer(scene)
        np.savez_compressed(a.aov, **r.aovs(w, h))
        r.close()
    im = render(scene, n_rays, a.n_bounces, seed=a.seed, width=w, height=h, flags=FLAG_PHYSICAL if a.physical else 0,
                scene_tonemap=a.scene_tonemap)
    out = a.out if a.out is not None else (scene.output if a.use_output else None)
    if out is not None:
        im.save(out, format="PPM") if out.lower().endswith(".pnm") else im.save(out)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
