"""Small pass over every pipeline of the library - the fused plane kernels, the LBVH wavefront with re-ordering and plane
culling, FP64 mode, records, physical mode, the closest-hit entry points, export / import, tone mapping - written to be run
under compute-sanitizer.  The tool is closed on this GPU pool (the wrapper answers "compute-sanitizer is closed on this pool
and stays closed"), so there is no sanitizer log in profiles/; the script still runs as a plain smoke pass."""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth
import torch

r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
acc, rec = r.trace_paths(2, 3, seed=3, width=64, height=64, records=True)
r.image(acc, 64, 64)
acc64, _ = r.trace_paths(1, 2, seed=3, precision=ptb.PREC_FP64)
r.image(acc64)
rays = np.random.default_rng(1).normal(size=(5000, 6)); rays[:, 2] -= 25.0
r.trace_closest(rays, ptb.PREC_FP64); r.trace_closest(rays.astype(np.float32))
t_rays = torch.as_tensor(rays, device="cuda:0").float().contiguous()
tri = torch.empty(5000, dtype=torch.int32, device="cuda:0"); t = torch.empty(5000, dtype=torch.float32, device="cuda:0")
r.trace_closest_noalloc(t_rays, tri, t)
r.trace_paths(2, 3, seed=3, width=64, height=64, flags=ptb.FLAG_SEPARATE_KERNELS)
r.trace_paths(2, 4, seed=3, width=64, height=64, flags=ptb.FLAG_PHYSICAL)
r.close()

scene = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8)
rl = ptb.Renderer(scene, device=0, accel=ptb.ACCEL_LBVH)
rl.trace_primary(64, 64)
rl.trace_paths(2, 3, seed=3, width=64, height=64, records=True)
rl.trace_paths(1, 3, seed=3, width=64, height=64, flags=ptb.FLAG_COUNT_NODES)
rl.trace_paths(1, 2, seed=3, width=32, height=32, precision=ptb.PREC_FP64)
rl.trace_paths(2, 4, seed=3, width=64, height=64, flags=ptb.FLAG_PHYSICAL)
blob = rl.export_bvh()
r2 = ptb.Renderer(scene, device=0, bvh=blob)
r2.trace_paths(1, 2, seed=3, width=32, height=32)
r2.close(); rl.close()
torch.cuda.synchronize()
print("sanitize_run: done")
