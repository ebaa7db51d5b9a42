import sys, os, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth, _lib
sc = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), int(sys.argv[1]) if len(sys.argv) > 1 else 177)
r = ptb.Renderer(sc, device=0)
W = H = 512
tri, t = r.trace_primary(W, H)
tri = tri.cpu().numpy()
tobj = np.asarray(r.flat["tri_obj"])
obj = np.where(tri >= 0, tobj[np.maximum(tri, 0)], -1)
names = ["left", "right", "floor", "back", "ceil", "small", "tall", "light"]
for k in range(8):
    px = np.flatnonzero(obj == k).astype(np.int32)
    if len(px) == 0: continue
    r.counters(reset=True)
    r.trace_paths(4, 1, seed=1, width=W, height=H, pixel_list=px, flags=_lib.FLAG_COUNT_NODES)
    c = r.counters(reset=True)
    print(names[k], len(px), "shadow rays", c["shadow_rays"], "nodes/ray %.1f tris/ray %.2f" % (c["nodes_visited_shadow"] / max(1, c["shadow_rays"]), c["tris_tested_shadow"] / max(1, c["shadow_rays"])))
for b in (1, 2, 3):
    r.counters(reset=True)
    r.trace_paths(2, b, seed=1, width=W, height=H, flags=_lib.FLAG_COUNT_NODES)
    c = r.counters(reset=True)
    print("bounces", b, "shadow nodes/ray %.1f closest nodes/ray %.1f" % (c["nodes_visited_shadow"] / max(1, c["shadow_rays"]), c["nodes_visited"] / max(1, c["closest_rays"])))
