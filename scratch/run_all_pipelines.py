"""Small end-to-end runs of every pipeline (both back ends, FP32 / FP64, fused / separate, physical mode): a quick
GPU smoke run outside pytest."""
import sys
sys.path.insert(0, "."); sys.path.insert(0, "tests")
import numpy as np
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth
from test_gpu_soup import soup_scene

cornell = ptb.Scene(ptb.CORNELL_SDL)
r = ptb.Renderer(cornell, device=0)
r.trace_primary(48, 40)
r.trace_paths(3, 3, seed=1, width=48, height=40)
r.trace_paths(2, 2, seed=1, width=32, height=32, flags=ptb.FLAG_SEPARATE_KERNELS, records=True)
r.trace_paths(2, 3, seed=1, width=32, height=32, flags=ptb.FLAG_PHYSICAL)
r.trace_paths(1, 2, seed=1, width=16, height=16, precision=ptb.PREC_FP64)
img = r.render(2, 2, width=32, height=32) if hasattr(r, "render") else None
r.close()
t = synth.tessellate(cornell, 6)                     # 1152 triangles -> LBVH, sort, bucketed light pick
r = ptb.Renderer(t, device=0)
r.trace_primary(40, 48)
r.trace_paths(3, 4, seed=2, width=40, height=48)
r.trace_paths(2, 3, seed=2, width=32, height=32, flags=ptb.FLAG_PHYSICAL)
r.trace_paths(1, 2, seed=2, width=16, height=16, precision=ptb.PREC_FP64)
r.trace_paths(1, 2, seed=2, width=16, height=16, flags=ptb.FLAG_COUNT_NODES)
blob = r.export_bvh()                                # plane ids travel with the hierarchy
r2 = ptb.Renderer(t, device=0, bvh=blob)
r2.trace_paths(2, 3, seed=2, width=24, height=24)
r2.image_scene(r2.trace_paths(1, 1, width=24, height=24)[0], 24, 24)
r2.close()
r.close()
with np.errstate(invalid="ignore", divide="ignore"):
    s = soup_scene(400, 3)
    s.flatten()
for accel in (ptb.ACCEL_BRUTE, ptb.ACCEL_LBVH):
    r = ptb.Renderer(s, device=0, accel=accel)
    r.trace_primary(32, 32)
    r.trace_paths(2, 3, seed=3, width=32, height=32)
    r.close()
ptb.release_workspace()
print("done")
