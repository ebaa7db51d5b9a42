"""Command line with the reference's flags (main.py:125-139):

    python -m pathtracerpython_b200 SCENE [--out F] [-r N] [-b N]

`--show-img` / `--show-scene` / `--show-normals` / `--show-screen` / `--show-inter` are accepted and
ignored: they open windows (PIL viewer, OpenGL debug viewer) and there is no display or graphics stack
on this box.  Extras (off by default, the reference ignores these scene keywords, A-Q19):
`--use-npaths` takes -r from the scene's `npaths`, `--seed` overrides the scene's `seed` as Philox key,
`--size W H` overrides `size`."""
import argparse

from . import render
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
    a = ap.parse_args(argv)
    scene = Scene(a.scene, quiet=False)
    print(f"Number of objects: {len(scene.objects)}")
    print(f"Number of triangles: {sum(len(o['geometry'].triangles) for o in scene.objects)}")
    n_rays = scene.npaths if (a.use_npaths and scene.npaths) else a.n_rays
    w, h = a.size if a.size else (None, None)
    im = render(scene, n_rays, a.n_bounces, seed=a.seed, width=w, height=h)
    if a.out is not None:
        im.save(a.out)
    return 0


if __name__ == "__main__":
    raise SystemExit(main())
