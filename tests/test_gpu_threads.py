"""include/ptb.h promises that any host thread may call in on any device and stream.  One process: two host threads on
two streams of one device sharing a scene handle; and, when the box has them, two devices - each with a 512-plane soup
(more than 48 KB of dynamic shared memory: a per-DEVICE kernel attribute) and a re-ordered LBVH frame (the sort's
64 KB opt-in, also per device).  Results must be bitwise equal to the single-threaded run on device 0."""
import threading

import numpy as np
import pytest

import pathtracerpython_b200 as ptb
from pathtracerpython_b200 import synth

pytestmark = pytest.mark.gpu


def _soup512():
    from test_gpu_soup import soup_scene
    from pathtracerpython_b200.scene import Obj
    sc = soup_scene(509, 14)
    geo = sc.objects[0]["geometry"]
    v = geo.vertexes.copy()
    v[0:18] += np.random.default_rng(1).normal(0.0, 0.2, (18, 3))
    sc.objects[0]["geometry"] = Obj.from_arrays(v, geo.faces)
    sc.flatten(refresh=True)
    return sc


def _frames(device, soup, tess, seeds):
    """Reference results, single-threaded on the default stream of `device`."""
    import torch
    with torch.cuda.device(device):
        rs = ptb.Renderer(soup, device=device, accel=ptb.ACCEL_BRUTE)
        rl = ptb.Renderer(tess, device=device, accel=ptb.ACCEL_LBVH)
        out = {}
        for sd in seeds:
            out[("soup", sd)] = rs.trace_paths(3, 3, seed=sd, width=64, height=64)[0].cpu()
            out[("lbvh", sd)] = rl.trace_paths(3, 4, seed=sd, width=96, height=96)[0].cpu()
        rs.close(); rl.close()
    return out


def test_two_host_threads_two_streams_one_device():
    import torch
    soup, tess = _soup512(), synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8)
    seeds = [3, 4, 5, 6]
    want = _frames(0, soup, tess, seeds)
    rs = ptb.Renderer(soup, device=0, accel=ptb.ACCEL_BRUTE)            # handles shared by both threads
    rl = ptb.Renderer(tess, device=0, accel=ptb.ACCEL_LBVH)
    got, errors = {}, []

    def work(my_seeds):
        try:
            st = torch.cuda.Stream(device=0)
            with torch.cuda.device(0), torch.cuda.stream(st):
                for _ in range(3):                                           # keep both threads launching for a while
                    for sd in my_seeds:
                        a = rs.trace_paths(3, 3, seed=sd, width=64, height=64)[0]
                        b = rl.trace_paths(3, 4, seed=sd, width=96, height=96)[0]
                        st.synchronize()
                        got[("soup", sd)], got[("lbvh", sd)] = a.cpu(), b.cpu()
        except Exception as e:      # noqa: BLE001
            errors.append(e)
    ts = [threading.Thread(target=work, args=(seeds[:2],)), threading.Thread(target=work, args=(seeds[2:],))]
    for t in ts:
        t.start()
    for t in ts:
        t.join()
    assert not errors, errors
    for k, v in want.items():
        assert torch.equal(v, got[k]), k
    rs.close(); rl.close()
    ptb.release_workspace()


def test_two_devices_in_one_process():
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs >= 2 GPUs")
    soup, tess = _soup512(), synth.tessellate(ptb.Scene(ptb.CORNELL_SDL), 8)
    a = _frames(0, soup, tess, [7])
    b = _frames(1, soup, tess, [7])            # second device of the process: needs its own kernel attributes
    for k in a:
        assert torch.equal(a[k], b[k]), k
    with torch.cuda.device(1):
        ptb.release_workspace(1)
