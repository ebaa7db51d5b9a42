"""Brute (plane list in shared memory) vs LBVH on random soups of growing size: where should ACCEL_AUTO switch?"""
import sys, time
sys.path.insert(0, "tests"); sys.path.insert(0, ".")
import numpy as np, torch
import pathtracerpython_b200 as ptb
from test_gpu_soup import soup_scene

for n in (32, 64, 128, 256, 509):
    sc = soup_scene(n, 21, n_obj=4)
    with np.errstate(invalid="ignore", divide="ignore"):
        sc.flatten()
    out = []
    for accel in (ptb.ACCEL_BRUTE, ptb.ACCEL_LBVH):
        r = ptb.Renderer(sc, device=0, accel=accel)
        r.trace_paths(16, 4, seed=1, width=512, height=512)
        torch.cuda.synchronize()
        t = time.perf_counter()
        r.trace_paths(64, 4, seed=1, width=512, height=512)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t
        out.append(512 * 512 * 64 / dt / 1e6)
        r.close()
    print(n, "triangles: brute %.1f  lbvh %.1f Msamples/s" % tuple(out), flush=True)
