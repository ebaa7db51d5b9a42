"""ptb_trace_closest_noalloc (include/ptb.h): the closest-hit query for callers that cannot allocate.  It has no
counterpart in the reference; its contract is "the same answer as ptb_trace_closest, with the scratch in a fixed device
array".  Through the C ABI (ctypes), two calls in a row - the array keeps the first call's contents, the second call is
shorter so stale rays lie right behind its own - against the allocating entry point, bit for bit (the tolerance of
the other closest-hit tests for two runs of the same kernels: equality), FP32 and FP64, plane-list and LBVH scenes.

Named test_zz_* so that pytest collects it last.  It holds alone in a fresh process
and after the other tests (nothing here depends on, or leaves, library state other than the scratch contents).
"""
import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import _lib, synth

pytestmark = pytest.mark.gpu


def _rays(n, seed):
    rng = np.random.default_rng(seed)
    o = np.stack([rng.uniform(-3.5, 3.5, n), rng.uniform(-3.5, 3.5, n), rng.uniform(-32, -17, n)], 1)
    return np.concatenate([o, rng.normal(size=(n, 3))], 1)


def _both(r, rays_np, precision, check_mem=False):
    import torch
    dt = torch.float64 if precision == ptb.PREC_FP64 else torch.float32
    rays = torch.as_tensor(rays_np, device=r.device).to(dt).contiguous()
    n = rays.shape[0]
    tri = torch.full((n,), -7, dtype=torch.int32, device=r.device)      # caller-owned outputs, made before the call
    t = torch.full((n,), -7.0, dtype=dt, device=r.device)
    want_tri, want_t = r.trace_closest(rays_np, precision)             # the existing entry point (cudaMallocAsync scratch)
    free_before = torch.cuda.mem_get_info(r.device)[0]
    r.trace_closest_noalloc(rays, tri, t, precision)
    torch.cuda.synchronize(r.device)
    if check_mem:                                                      # once everything is warm (lazily loaded kernels take memory)
        assert torch.cuda.mem_get_info(r.device)[0] >= free_before     # the call took no device memory
    return tri.cpu().numpy(), t.cpu().numpy(), want_tri.cpu().numpy(), want_t.cpu().numpy()


@pytest.mark.parametrize("accel", ["planes", "lbvh"])
def test_noalloc_equals_allocating_entry_twice_in_a_row(accel):
    if accel == "planes":
        r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    else:
        r = ptb.Renderer(synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8), device=0, accel=ptb.ACCEL_LBVH)     # 2048 triangles
    for call, (n, seed) in enumerate([(30000, 21), (17001, 22)]):        # second call shorter: stale queue entries behind it
        for precision in (ptb.PREC_FP32, ptb.PREC_FP64):
            tri, t, want_tri, want_t = _both(r, _rays(n, seed), precision, check_mem=call == 1)
            assert (want_tri >= 0).mean() > 0.99                         # closed room: a line always hits something
            assert np.array_equal(tri, want_tri), f"call {call}: triangle ids differ"
            assert np.array_equal(t, want_t), f"call {call}: ray parameters differ"


def test_noalloc_limit_and_arguments():
    import torch
    r = ptb.Renderer(ptb.CORNELL_SDL, device=0)
    L = _lib.lib()
    cap = _lib.TRACE_CLOSEST_NOALLOC_MAX_RAYS
    assert cap == 131072                  # include/ptb.h PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS
    n = cap + 1
    rays = torch.zeros((n, 6), dtype=torch.float32, device=r.device)
    rays[:, 5] = 1.0
    tri = torch.empty(n, dtype=torch.int32, device=r.device)
    t = torch.empty(n, dtype=torch.float32, device=r.device)
    with pytest.raises(ptb.PtbError, match="at most 131072"):
        r.trace_closest_noalloc(rays, tri, t)
    # exactly the maximum is served, and equals the allocating entry point
    rays_np = _rays(cap, 23).astype(np.float32)
    got_tri, got_t, want_tri, want_t = _both(r, rays_np, ptb.PREC_FP32)
    assert np.array_equal(got_tri, want_tri) and np.array_equal(got_t, want_t)
    # n == 0 is a no-op, null pointers are refused
    assert L.ptb_trace_closest_noalloc(r._handle, rays.data_ptr(), 0, ptb.PREC_FP32, tri.data_ptr(), t.data_ptr(), None) == 0
    assert L.ptb_trace_closest_noalloc(r._handle, None, 4, ptb.PREC_FP32, tri.data_ptr(), t.data_ptr(), None) == -1
