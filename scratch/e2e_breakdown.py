"""Where the end-to-end step of an LBVH scene spends its time beyond the frame: Renderer creation pieces."""
import sys, os, time, numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth
n = int(sys.argv[1]) if len(sys.argv) > 1 else 177
sc = synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), n)
sc.flatten(pin=True)
def t(): torch.cuda.synchronize(); return time.perf_counter()
for it in range(3):
    a = t(); r = ptb.Renderer(sc, device=0); b = t()
    r.trace_paths(1, 5, seed=1, width=1024, height=1024); c = t()
    r.trace_paths(1, 5, seed=1, width=1024, height=1024); d = t()
    img = r.image(torch.zeros((1024 * 1024, 4), device="cuda")); e = t()
    r.close(); del r; f = t()
    print("iter", it, "create %.1f ms, first frame(1spp) %.1f ms, second %.1f ms, image %.1f ms, close %.1f ms" % (1e3*(b-a), 1e3*(c-b), 1e3*(d-c), 1e3*(e-d), 1e3*(f-e)))
import cProfile, pstats
pr = cProfile.Profile(); pr.enable(); r = ptb.Renderer(sc, device=0); torch.cuda.synchronize(); pr.disable()
pstats.Stats(pr).sort_stats("cumulative").print_stats(12)
print("---- bench-like e2e step at 64 spp")
from pathtracerpython_b200 import dist as pdist
W = H = 1024
full = torch.zeros((W * H, 4), dtype=torch.float32, device="cuda")
host_img = torch.empty((H, W, 3), dtype=torch.uint8).pin_memory()
r0 = ptb.Renderer(sc, device=0)
for it in range(3):
    a = t(); img, _ = pdist.render_step(r0, 64, 5, 9, W, H, 0, 1, "tile", H, full); b = t()
    print("resident frame %.1f ms" % (1e3 * (b - a)))
for it in range(3):
    a = t(); rr = ptb.Renderer(sc, device=0); b = t()
    img, _ = pdist.render_step(rr, 64, 5, 9, W, H, 0, 1, "tile", H, full); c = t()
    host_img.copy_(img, non_blocking=True); torch.cuda.current_stream().synchronize(); d = t()
    rr.close(); e = t()
    print("e2e: create %.1f, frame %.1f, image copy %.1f, close %.1f ms" % (1e3*(b-a), 1e3*(c-b), 1e3*(d-c), 1e3*(e-d)))
