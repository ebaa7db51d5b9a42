import sys, os, struct, numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth, _lib
n = int(sys.argv[1]) if len(sys.argv) > 1 else 60
sc = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), n)
r = ptb.Renderer(sc, device=0, accel=ptb.ACCEL_LBVH)
blob = r.export_bvh().cpu().numpy()
magic, ver, n_tri, n_obj_tri, has_occ, has_pid, pad = struct.unpack("<IIiiiif", blob[:28].tobytes())
off = struct.unpack("<8Q", blob[32:96].tobytes())
print("version", ver, "n_tri", n_tri, "has_pid", has_pid, "offsets", off)
pid = blob[off[6]:off[6] + n_tri]
tobj = np.asarray(r.flat["tri_obj"])
for o in range(8):
    u, c = np.unique(pid[tobj == o], return_counts=True)
    print("object", o, dict(zip(u.tolist(), c.tolist())))
nodes = blob[off[0]:off[0] + 64 * (n_tri - 1)].view(np.int32).reshape(-1, 16)
meta = nodes[:, 13]
p0, p1 = meta & 0xff, (meta >> 8) & 0xff
print("children with a plane id: %.3f" % ((p0 > 0).mean() / 2 + (p1 > 0).mean() / 2))
W = 256
tri, _ = r.trace_primary(W, W)
tri = tri.cpu().numpy()
obj = np.where(tri >= 0, tobj[np.maximum(tri, 0)], -1)
for k in (4, 5, 6):
    px = np.flatnonzero(obj == k).astype(np.int32)
    for env in (0, 1):
        r.counters(reset=True)
        r.trace_paths(4, 1, seed=1, width=W, height=W, pixel_list=px, flags=_lib.FLAG_COUNT_NODES)
        c = r.counters(reset=True)
        print("object", k, "pixels", len(px), "shadow nodes/ray %.2f tris %.2f" % (c["nodes_visited_shadow"] / max(1, c["shadow_rays"]), c["tris_tested_shadow"] / max(1, c["shadow_rays"])))
        break
    # which faces do the pixels see
    u, c = np.unique(pid[tri[px]], return_counts=True)
    print("   planes seen:", dict(zip(u.tolist(), c.tolist())))
