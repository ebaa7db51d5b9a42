"""Host side of the hot path: the replacement for main.py's sample/bounce loops (main.py:165-293).

`render(scene_or_path, n_rays, n_bounces)` -> PIL.Image is the drop-in for `main()`'s body: same
inputs (`Scene`, `-r`, `-b`), same output (`make_image`'s 8-bit RGB image).  `Renderer` is the
lower-level object: it owns the device scene handle (libptb.so) and the torch tensors used as
workspace / accumulation buffers, and exposes the three parity seams:

    trace_primary()  <-> one round of intersect_objects over the primary rays  (main.py:199-204)
    trace_paths()    <-> the sample loop x bounce loop                         (main.py:186-271)
    image()          <-> make_image                                           (utils.py:150-161)

PyTorch is used for device memory and streams only; all arithmetic happens in libptb.so.  There is
no CPU fallback: without CUDA or without the library these calls raise.
"""
import ctypes as C

import numpy as np

from . import _lib
from ._lib import PtbError
from .scene import Scene


def _torch():
    import torch
    if not torch.cuda.is_available():
        raise PtbError("CUDA device required: pathtracerpython_b200 has no CPU fallback")
    return torch


# One workspace per device, shared by every Renderer of the process: the wavefront state of a batch is tens of GB, and
# allocating it again for each new Renderer (a scene per frame in an animation, bench.py's end-to-end step) costs more
# than building the scene.  Keyed by (device, CUDA stream): calls on one stream are ordered, so they can share it; a second
# stream gets its own.  It only grows; release_workspace() hands it back.
_WORKSPACES = {}


def release_workspace(device=None):
    """Free the shared wavefront workspace of `device` (all devices when None)."""
    for k in [k for k in _WORKSPACES if device is None or k[0] == _dev_key(device)]:
        del _WORKSPACES[k]
    _lib.check(_lib.lib().ptb_pool_trim(), "ptb_pool_trim")      # and the scene pool's unused memory (current device)


def _dev_key(device):
    import torch
    d = torch.device("cuda", device) if isinstance(device, int) else torch.device(device)
    return d.index if d.index is not None else torch.cuda.current_device()


class Renderer:
    def __init__(self, scene, device=None, accel=_lib.ACCEL_AUTO, workspace_bytes=None, bvh=None, device_arrays=None):
        """`bvh`: a uint8 device tensor from another Renderer's `export_bvh()` of the SAME scene (possibly
        received from another rank, dist.replicate_renderer): the hierarchies are copied from it
        (ptb_bvh_import) instead of being built.
        `device_arrays`: (verts float64 [n,3,3], normals float64 [n,3], tri_obj int32 [n]) torch tensors already on the
        renderer's device (PTB_SCENE_ARRAYS_ON_DEVICE; dist.gather_scene_arrays assembles them from per-rank slices
        over NVLink): the big arrays are then not read from the host at all.  LBVH scenes only."""
        torch = _torch()
        self.torch = torch
        self.scene = Scene(scene) if isinstance(scene, str) else scene
        self.device = torch.device("cuda", torch.cuda.current_device() if device is None else device)
        self.flat = self.scene.flatten() if hasattr(self.scene, "flatten") else self.scene
        self.width, self.height = int(self.flat["width"]), int(self.flat["height"])
        self.workspace_bytes = workspace_bytes
        self._ws = None
        self._handle = C.c_void_p()
        f = self.flat
        self._keep = [np.ascontiguousarray(f[k], dt) for k, dt in
                      (("verts", np.float64), ("normals", np.float64), ("tri_obj", np.int32), ("mat", np.float64),
                       ("light_areas", np.float64))]
        d = _lib.SceneDesc()
        d.n_tri = self._keep[0].shape[0]
        d.n_tri_obj = int(f["n_tri_obj"])
        d.n_obj = self._keep[3].shape[0]
        d.n_light_tri = d.n_tri - d.n_tri_obj
        d.verts, d.normals, d.tri_obj, d.mat, d.light_areas = [a.ctypes.data for a in self._keep]
        if device_arrays is not None:
            dv, dn, do = device_arrays
            if (dv.dtype != torch.float64 or dn.dtype != torch.float64 or do.dtype != torch.int32 or dv.numel() != 9 * d.n_tri
                    or dn.numel() != 3 * d.n_tri or do.numel() != d.n_tri or any(t.device != self.device or not t.is_contiguous() for t in (dv, dn, do))):
                raise PtbError("device_arrays: contiguous float64 [n,3,3], float64 [n,3], int32 [n] tensors on the renderer's device")
            d.verts, d.normals, d.tri_obj = dv.data_ptr(), dn.data_ptr(), do.data_ptr()
            d.reserved = _lib.SCENE_ARRAYS_ON_DEVICE
        d.eye[:] = [float(x) for x in f["eye"]]
        d.ortho[:] = [float(x) for x in f["ortho"]]
        d.ambient = float(f["ambient"])
        d.light_color[:] = [float(x) for x in f["light_color"][:3]]
        d.accel = accel
        with torch.cuda.device(self.device):
            if bvh is None:
                _lib.check(_lib.lib().ptb_scene_create(C.byref(d), self._stream(), C.byref(self._handle)), "ptb_scene_create")
            else:
                if bvh.dtype != torch.uint8 or bvh.device != self.device or not bvh.is_contiguous():
                    raise PtbError("bvh must be a contiguous uint8 tensor on the renderer's device")
                _lib.check(_lib.lib().ptb_bvh_import(C.byref(d), bvh.data_ptr(), bvh.numel(), self._stream(),
                                                     C.byref(self._handle)), "ptb_bvh_import")
        self.n_tri = d.n_tri

    # ---- plumbing
    def _stream(self):
        return C.c_void_p(self.torch.cuda.current_stream(self.device).cuda_stream)

    def close(self):
        if getattr(self, "_handle", None) and self._handle.value:
            _lib.lib().ptb_scene_destroy(self._handle)
            self._handle = C.c_void_p()

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def pixels(self, width=None, height=None, shard=0, n_shards=1, tile=4096, pixel_list=None):
        px = _lib.Pixels()
        px.width = int(width or self.width)
        px.height = int(height or self.height)
        px.shard, px.n_shards, px.tile = int(shard), int(n_shards), int(tile)
        keep = None
        if pixel_list is not None:
            lst = np.asarray(pixel_list, np.int32).reshape(-1)
            px.n_pixels = int(lst.size)
            # an empty list is still a list (NULL would mean "the whole shard"): keep a 1-element buffer behind it
            keep = self.torch.as_tensor(lst if lst.size else np.zeros(1, np.int32), device=self.device)
            px.d_pixels = keep.data_ptr()
        return px, keep

    def pixel_count(self, px):
        return int(_lib.lib().ptb_pixel_count(C.byref(px)))

    def workspace(self, nbytes):
        key = (_dev_key(self.device), self.torch.cuda.current_stream(self.device).cuda_stream)
        ws = _WORKSPACES.get(key)
        if ws is None or ws.numel() < nbytes:
            _WORKSPACES.pop(key, None)
            ws = None
            ws = _WORKSPACES[key] = self.torch.empty(int(nbytes), dtype=self.torch.uint8, device=self.device)
        self._ws = ws
        return ws

    # ---- parity seam 1
    def trace_primary(self, width=None, height=None, pixel_list=None, precision=_lib.PREC_FP32, flags=0):
        """Closest-hit triangle id (-1 = miss) and ray parameter for the primary rays."""
        torch = self.torch
        px, keep = self.pixels(width, height, pixel_list=pixel_list)
        n = self.pixel_count(px)
        tri = torch.empty(n, dtype=torch.int32, device=self.device)
        t = torch.empty(n, dtype=torch.float32, device=self.device)
        ws = self.workspace(int(_lib.lib().ptb_trace_primary_workspace(n, precision)))
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_trace_primary(self._handle, C.byref(px), precision, flags, tri.data_ptr(), t.data_ptr(),
                                                    ws.data_ptr(), ws.numel(), self._stream()), "ptb_trace_primary")
        return tri, t

    def aovs(self, width=None, height=None):
        """Primary-hit buffers as numpy arrays in ray-index order (r = ix * H + iy): the data behind the
        reference's debug views (`--show-inter`, `--show-normals`; plot.py:28-42, 98-105), headless."""
        W, H = int(width or self.width), int(height or self.height)
        tri, t = self.trace_primary(W, H)
        tri, t = tri.cpu().numpy(), t.cpu().numpy().astype(np.float64)
        f = self.flat
        eye, (x0, y0, x1, y1) = np.asarray(f["eye"], np.float64), [float(v) for v in f["ortho"]]
        xs, ys = np.linspace(x0, x1, W), np.linspace(y0, y1, H)             # utils.py:64-69
        d = np.stack([np.repeat(xs, H), np.tile(ys, W), np.zeros(W * H)], 1) - eye
        d /= np.linalg.norm(d, axis=1, keepdims=True)
        hit = tri >= 0
        safe = np.where(hit, tri, 0)
        return {"tri": tri, "t": np.where(hit, t, np.nan), "point": np.where(hit[:, None], eye + d * t[:, None], np.nan),
                "normal": np.where(hit[:, None], np.asarray(f["normals"])[safe], np.nan),
                "object": np.where(hit, np.asarray(f["tri_obj"])[safe], -1), "width": W, "height": H}

    def trace_closest(self, rays, precision=_lib.PREC_FP32):
        """intersect_objects for caller-supplied rays [n, 6] (origin, direction)."""
        torch = self.torch
        dt = torch.float64 if precision == _lib.PREC_FP64 else torch.float32
        rays = torch.as_tensor(np.asarray(rays), device=self.device).to(dt).contiguous().reshape(-1, 6)
        n = rays.shape[0]
        tri = torch.empty(n, dtype=torch.int32, device=self.device)
        t = torch.empty(n, dtype=dt, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_trace_closest(self._handle, rays.data_ptr(), n, precision, tri.data_ptr(), t.data_ptr(),
                                                    self._stream()), "ptb_trace_closest")
        return tri, t

    def trace_closest_noalloc(self, rays, tri, t, precision=_lib.PREC_FP32):
        """ptb_trace_closest_noalloc: the same query into caller-owned DEVICE tensors (`rays` [n, 6] float32 / float64
        contiguous, `tri` int32 [n], `t` like rays [n]); the library allocates nothing (its scratch is a fixed device
        array), at most `_lib.TRACE_CLOSEST_NOALLOC_MAX_RAYS` rays per call, calls must not overlap on the GPU."""
        n = rays.shape[0]
        with self.torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_trace_closest_noalloc(self._handle, rays.data_ptr(), n, precision, tri.data_ptr(),
                                                            t.data_ptr(), self._stream()), "ptb_trace_closest_noalloc")
        return tri, t

    # ---- parity seam 2 / the hot path
    def trace_paths(self, n_rays, n_bounces, seed=0, width=None, height=None, shard=0, n_shards=1, tile=4096,
                    pixel_list=None, sample_begin=0, precision=_lib.PREC_FP32, flags=0, records=False, accum=None,
                    samples_per_batch=0):
        """Adds samples [sample_begin, sample_begin + n_rays) of every pixel of the set to `accum`
        ([n_pixel_slots, 4] float32: r, g, b sums and sample count).  Returns (accum, records | None)."""
        torch = self.torch
        p = _lib.RenderParams()
        p.px, keep = self.pixels(width, height, shard, n_shards, tile, pixel_list)
        npix = self.pixel_count(p.px)
        p.sample_begin, p.sample_end = int(sample_begin), int(sample_begin + n_rays)
        p.bounces, p.seed, p.precision, p.flags = int(n_bounces), int(seed) & 0xFFFFFFFF, precision, flags
        p.samples_per_batch = int(samples_per_batch)
        if accum is None:
            accum = torch.zeros((npix, 4), dtype=torch.float32, device=self.device)
        if (pixel_list is not None and npix == 0) or n_rays <= 0:   # empty pixel list / sample range: nothing to launch
            return accum, None
        rec, rec_t = None, None
        if records:
            sh = (npix, n_rays, n_bounces)
            rec_t = {"tri": torch.empty(sh, dtype=torch.int32, device=self.device),
                     "contrib": torch.empty(sh + (3,), dtype=torch.float32, device=self.device),
                     "k_after": torch.empty(sh, dtype=torch.float32, device=self.device),
                     "lobe": torch.empty(sh, dtype=torch.int8, device=self.device),
                     "occ": torch.empty(sh, dtype=torch.int8, device=self.device),
                     "leak": torch.empty(sh, dtype=torch.int8, device=self.device)}
            if precision == _lib.PREC_FP64:
                rec_t["contrib64"] = torch.empty(sh + (3,), dtype=torch.float64, device=self.device)
                rec_t["k_after64"] = torch.empty(sh, dtype=torch.float64, device=self.device)
            rec = _lib.PathRecords()
            for k, v in rec_t.items():
                setattr(rec, "d_" + k, v.data_ptr())
        L = _lib.lib()
        if self.workspace_bytes:
            ws_bytes = self.workspace_bytes
        else:
            # default: as many samples per batch as fit in ~1/4 of free memory, at most all of them
            # (what the shared workspace already holds counts as free, so the batch size does not drift from call to call)
            free, _ = torch.cuda.mem_get_info(self.device)
            held = _WORKSPACES.get((_dev_key(self.device), torch.cuda.current_stream(self.device).cuda_stream))
            free += held.numel() if held is not None else 0
            one = L.ptb_trace_paths_workspace(C.byref(p), 1)
            spb = max(1, min(n_rays, int((free // 4) // max(one, 1)), (2 ** 31 - 1) // max(npix, 1)))
            if samples_per_batch:
                spb = min(spb, samples_per_batch)
            ws_bytes = L.ptb_trace_paths_workspace(C.byref(p), spb)
            if held is not None and ws_bytes > held.numel():
                # Free memory moves by a few MB between calls (a scene created, an image buffer), and one sample per batch
                # more than last time would throw a workspace of tens of GB away and allocate a larger one - 20 ms of
                # cudaFree / cudaMalloc.  If what is held takes at least 3/4
                # of the batch just computed, the batch is cut to what is held.
                fit = int(min(spb, held.numel() // max(one, 1)))
                while fit > 0 and L.ptb_trace_paths_workspace(C.byref(p), fit) > held.numel():
                    fit -= 1
                if 4 * fit >= 3 * spb:
                    spb, ws_bytes = fit, L.ptb_trace_paths_workspace(C.byref(p), fit)
        ws = self.workspace(ws_bytes)
        with torch.cuda.device(self.device):
            # the size asked for, not the (shared, possibly larger) tensor's: a caller-imposed budget stays a budget
            _lib.check(L.ptb_trace_paths(self._handle, C.byref(p), ws.data_ptr(), int(ws_bytes), accum.data_ptr(),
                                         C.byref(rec) if rec is not None else None, self._stream()), "ptb_trace_paths")
        return accum, rec_t

    # ---- output seam
    def image(self, accum_full, width=None, height=None):
        """make_image on a full-frame accumulation buffer (ray-index order) -> uint8 tensor [H, W, 3]."""
        torch = self.torch
        W, H = int(width or self.width), int(height or self.height)
        rgb = torch.empty((H, W, 3), dtype=torch.uint8, device=self.device)
        ws = self.workspace(256)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_tonemap(accum_full.data_ptr(), W, H, rgb.data_ptr(), ws.data_ptr(), self._stream()),
                       "ptb_tonemap")
        return rgb

    def image_scene(self, accum_full, width=None, height=None, tonemapping=None, background=None, primary_tri=None):
        """Opt-in alternative to `image`: the scene's `tonemapping` (v / (v + tm)) and `background` keywords, which the
        reference parses and ignores (scene_reader.py:158-159, 167-168) -> uint8 tensor [H, W, 3].  `primary_tri`
        (int32 per ray index, < 0 = miss) selects the background pixels; by default the primary rays are traced."""
        torch = self.torch
        W, H = int(width or self.width), int(height or self.height)
        tm = self.flat.get("tonemapping") if tonemapping is None else tonemapping
        bg = self.flat.get("background") if background is None else background
        tm = 0.0 if tm is None else float(tm)
        bg = (C.c_double * 3)(*[float(x) for x in (bg if bg is not None else (0.0, 0.0, 0.0))][:3])
        if primary_tri is None:
            primary_tri, _ = self.trace_primary(W, H)
        rgb = torch.empty((H, W, 3), dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_tonemap_scene(accum_full.data_ptr(), primary_tri.data_ptr(), W, H, tm, bg,
                                                    rgb.data_ptr(), self._stream()), "ptb_tonemap_scene")
        return rgb

    def scatter(self, px, accum_shard, accum_full):
        with self.torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_scatter_accum(C.byref(px), accum_shard.data_ptr(), accum_full.data_ptr(), self._stream()),
                       "ptb_scatter_accum")
        return accum_full

    def export_bvh(self, out=None):
        """Both hierarchies of an LBVH scene as ONE uint8 device tensor (ptb_bvh_export) that `Renderer(scene,
        bvh=...)` on any GPU turns back into a device scene without the build.  Raises for shared-memory scenes."""
        torch = self.torch
        L = _lib.lib()
        nbytes = int(L.ptb_bvh_export_bytes(self._handle))
        if nbytes == 0:
            raise PtbError("export_bvh: the scene has no LBVH (small scenes live in shared memory)")
        if out is None:
            out = torch.empty(nbytes, dtype=torch.uint8, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(L.ptb_bvh_export(self._handle, out.data_ptr(), out.numel(), self._stream()), "ptb_bvh_export")
        return out

    def info(self):
        """-> dict(accel, n_nodes, build_ms) of the device scene."""
        a, n, ms = C.c_int32(), C.c_int32(), C.c_double()
        _lib.check(_lib.lib().ptb_scene_info(self._handle, C.byref(a), C.byref(n), C.byref(ms)), "ptb_scene_info")
        return {"accel": {1: "brute", 2: "lbvh"}.get(a.value, a.value), "n_nodes": n.value, "build_ms": ms.value,
                "plane_culling": bool(_lib.lib().ptb_scene_plane_culling(self._handle))}

    def kat(self, what, data, precision=_lib.PREC_FP64, width=None, height=None, seed=0):
        """ptb_kat: known answers of device functions without a seam of their own ("rotate", "light", "ray_dir")."""
        torch = self.torch
        code = {"rotate": 0, "light": 1, "ray_dir": 2}[what]
        x = torch.as_tensor(np.ascontiguousarray(data, np.float64), device=self.device).contiguous()
        n = x.shape[0]
        out = torch.empty((n, 4 if code == 1 else 3), dtype=torch.float64, device=self.device)
        with torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_kat(self._handle, code, precision, x.data_ptr(), n, out.data_ptr(),
                                          int(width or self.width), int(height or self.height), int(seed) & 0xFFFFFFFF,
                                          self._stream()), "ptb_kat")
        return out.cpu().numpy()

    def counters(self, reset=False):
        c = _lib.Counters()
        with self.torch.cuda.device(self.device):
            _lib.check(_lib.lib().ptb_counters_read(self._handle, C.byref(c), self._stream()), "ptb_counters_read")
            if reset:
                _lib.check(_lib.lib().ptb_counters_reset(self._handle, self._stream()), "ptb_counters_reset")
        return {k: int(getattr(c, k)) for k, _ in _lib.Counters._fields_}


def render(scene, n_rays=1, n_bounces=1, seed=None, width=None, height=None, device=None, precision=_lib.PREC_FP32,
           flags=0, return_radiance=False, scene_tonemap=False):
    """Drop-in for the body of main() (main.py:165-293): scene (path or Scene) -> PIL.Image.
    `seed` defaults to the scene's `seed` keyword (parsed but unused by the reference, A-Q19).
    `scene_tonemap=True` (opt-in, not the reference's output) applies the scene's `tonemapping` / `background`
    keywords (Renderer.image_scene) instead of make_image's min/max normalisation."""
    from PIL import Image
    r = Renderer(scene, device=device)
    try:
        if seed is None:
            seed = r.flat.get("seed") or 0
        accum, _ = r.trace_paths(n_rays, n_bounces, seed=seed, width=width, height=height, precision=precision, flags=flags)
        rgb = r.image_scene(accum, width, height) if scene_tonemap else r.image(accum, width, height)
        im = Image.fromarray(rgb.cpu().numpy())
        if return_radiance:
            a = accum.cpu().numpy()
            return im, a[:, :3] / np.maximum(a[:, 3:4], 1)
        return im
    finally:
        r.close()


class ProgressiveFrame:
    """Progressive rendering of one frame as a CUDA graph (ptb_trace_paths_at): the launch sequence of `n_rays` samples per
    pixel - camera rays, the per-bounce kernels, the batch reduce, the tone map - is captured once; every `step()` replays it
    with the sample offset, a device word the graph itself advances, `n_rays` further on.  After k steps `accum` holds the same
    sums, bit for bit, as one `trace_paths(k * n_rays, ..., samples_per_batch=n_rays)`.  For frames that are launch-bound
    (the reference's shipped 40 x 40 scene: 0.17 ms of launches for 0.03 ms of kernels) a replay costs one graph launch."""

    def __init__(self, renderer, n_rays, n_bounces, seed=0, width=None, height=None, tonemap=True, flags=0):
        torch = renderer.torch
        self.r, self.n_rays = renderer, int(n_rays)
        L = _lib.lib()
        p = _lib.RenderParams()
        p.px, self._keep = renderer.pixels(width, height)
        self.width, self.height = int(p.px.width), int(p.px.height)
        npix = renderer.pixel_count(p.px)
        p.sample_begin, p.sample_end = 0, self.n_rays
        p.bounces, p.seed, p.precision, p.flags = int(n_bounces), int(seed) & 0xFFFFFFFF, _lib.PREC_FP32, int(flags)
        p.samples_per_batch = self.n_rays
        self._p = p
        dev = renderer.device
        self.accum = torch.zeros((npix, 4), dtype=torch.float32, device=dev)
        self.offset = torch.zeros(1, dtype=torch.int32, device=dev)
        self.rgb = torch.empty((self.height, self.width, 3), dtype=torch.uint8, device=dev) if tonemap else None
        self._ws_bytes = int(L.ptb_trace_paths_workspace(C.byref(p), self.n_rays))
        self._ws = torch.empty(self._ws_bytes, dtype=torch.uint8, device=dev)
        self._ws_tm = torch.empty(256, dtype=torch.uint8, device=dev)
        self.steps = 0
        with torch.cuda.device(dev):
            side = torch.cuda.Stream(device=dev)
            side.wait_stream(torch.cuda.current_stream(dev))
            with torch.cuda.stream(side):           # once outside the capture: module loading, kernel attributes
                self._launch()
            side.synchronize()
            self.accum.zero_(); self.offset.zero_()
            torch.cuda.current_stream(dev).synchronize()
            self.graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.graph):
                self._launch()

    def _launch(self):
        torch, r, L = self.r.torch, self.r, _lib.lib()
        st = C.c_void_p(torch.cuda.current_stream(r.device).cuda_stream)
        _lib.check(L.ptb_trace_paths_at(r._handle, C.byref(self._p), self._ws.data_ptr(), self._ws_bytes, self.accum.data_ptr(), None,
                                        self.offset.data_ptr(), st), "ptb_trace_paths_at")
        self.offset.add_(self.n_rays)
        if self.rgb is not None:
            _lib.check(L.ptb_tonemap(self.accum.data_ptr(), self.width, self.height, self.rgb.data_ptr(), self._ws_tm.data_ptr(), st), "ptb_tonemap")

    def step(self):
        """n_rays more samples per pixel -> (accum, rgb)."""
        self.graph.replay()
        self.steps += 1
        return self.accum, self.rgb

    def reset(self):
        self.accum.zero_(); self.offset.zero_(); self.steps = 0
