"""ProgressiveFrame (ptb_trace_paths_at): a frame's launch sequence captured once as a CUDA graph, the sample offset a device
word the graph advances itself.  k replays must hold, bit for bit, the sums of one eager call over k times the samples in
batches of the same size; the image is the eager image; a second graph on the same renderer does not disturb the first."""
import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("kind", ["planes", "lbvh"])
def test_graph_replays_equal_the_eager_call(kind):
    import torch
    if kind == "planes":
        r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
        W = H = 40
    else:
        r = ptb.Renderer(synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8), device=0, accel=ptb.ACCEL_LBVH)
        W = H = 64
    n, nb, k = 5, 3, 4
    pf = ptb.ProgressiveFrame(r, n, nb, seed=11, width=W, height=H)
    for _ in range(k):
        accum, rgb = pf.step()
    torch.cuda.synchronize()
    assert int(pf.offset.item()) == k * n
    want, _ = r.trace_paths(k * n, nb, seed=11, width=W, height=H, samples_per_batch=n)
    assert np.array_equal(accum.cpu().numpy(), want.cpu().numpy())
    assert np.array_equal(rgb.cpu().numpy(), r.image(want, W, H).cpu().numpy())
    # the first replay alone == the first n samples; reset starts over
    pf.reset()
    accum, _ = pf.step()
    first, _ = r.trace_paths(n, nb, seed=11, width=W, height=H)
    assert np.array_equal(accum.cpu().numpy(), first.cpu().numpy())
    # two graphs side by side
    pf2 = ptb.ProgressiveFrame(r, n, nb, seed=12, width=W, height=H, tonemap=False)
    pf2.step(); pf.step(); pf2.step()
    torch.cuda.synchronize()
    w1, _ = r.trace_paths(2 * n, nb, seed=11, width=W, height=H, samples_per_batch=n)
    w2, _ = r.trace_paths(2 * n, nb, seed=12, width=W, height=H, samples_per_batch=n)
    assert np.array_equal(pf.accum.cpu().numpy(), w1.cpu().numpy())
    assert np.array_equal(pf2.accum.cpu().numpy(), w2.cpu().numpy())
    r.close()
