"""Upper bound for object-local frames: the Cornell blocks rotated about the vertical axis until their faces are
axis-aligned (a different scene, for counting only), tessellated like config 4.  Prints nodes per ray and frame time
next to the original scene's."""
import sys, os, math, copy, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth, _lib
from pathtracerpython_b200.scene import Obj

def aligned(scene):
    out = copy.copy(scene); out.objects = []
    for k, o in enumerate(scene.objects):
        d = dict(o)
        t = np.asarray(o["geometry"].triangles, np.float64)
        if k in (5, 6):
            nrm = np.asarray(o["geometry"].normals, np.float64)
            side = nrm[np.argmax(np.abs(nrm[:, 0]) * (np.abs(nrm[:, 1]) < 1e-6))]
            ang = math.atan2(side[2], side[0])
            c, s = math.cos(ang), math.sin(ang)
            ctr = t.reshape(-1, 3).mean(0)
            p = t.reshape(-1, 3) - ctr
            q = np.stack([c * p[:, 0] + s * p[:, 2], p[:, 1], -s * p[:, 0] + c * p[:, 2]], 1) + ctr
            t = q.reshape(-1, 3, 3)
            print("object", k, "rotated by", math.degrees(ang))
        v = t.reshape(-1, 3)
        d["geometry"] = Obj.from_arrays(v, np.arange(len(v), dtype=np.int64).reshape(-1, 3), "object")
        out.objects.append(d)
    return out

def measure(sc, name, W=1024, spp=16, nb=5):
    r = ptb.Renderer(sc, device=0)
    r.trace_paths(2, nb, seed=1, width=W, height=W)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record(); r.trace_paths(spp, nb, seed=1, width=W, height=W); e1.record(); torch.cuda.synchronize()
    ms = e0.elapsed_time(e1)
    r.counters(reset=True)
    r.trace_paths(1, nb, seed=1, width=W, height=W, flags=_lib.FLAG_COUNT_NODES)
    c = r.counters(reset=True)
    print(name, "%.1f Msamples/s" % (W * W * spp / ms / 1e3), "shadow nodes/ray %.1f tris %.2f | closest nodes/ray %.1f tris %.2f" % (
        c["nodes_visited_shadow"] / c["shadow_rays"], c["tris_tested_shadow"] / c["shadow_rays"],
        c["nodes_visited"] / c["closest_rays"], c["tris_tested"] / c["closest_rays"]))
    r.close()

base = ptb.Scene(ptb.CORNELL_SDL)
n = int(sys.argv[1]) if len(sys.argv) > 1 else 177
measure(synth.tessellate(base, n), "original")
measure(synth.tessellate(aligned(base), n), "blocks axis-aligned")
