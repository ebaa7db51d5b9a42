"""ctypes binding of libptb.so (include/ptb.h).  There is no CPU fallback: if the CUDA library is
missing or a call fails, this module raises."""
import ctypes as C
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("PTB_LIB_PATH") or os.path.join(_HERE, "libptb.so")      # override: A/B builds of the same source

PREC_FP32, PREC_FP64 = 0, 1
ACCEL_AUTO, ACCEL_BRUTE, ACCEL_LBVH = 0, 1, 2
SCENE_ARRAYS_ON_DEVICE = 1
FLAG_NO_FP64_RETRACE, FLAG_SEPARATE_KERNELS, FLAG_COUNT_NODES, FLAG_PHYSICAL, FLAG_NO_ROULETTE = 1, 2, 4, 8, 16
TRACE_CLOSEST_NOALLOC_MAX_RAYS = 131072      # include/ptb.h PTB_TRACE_CLOSEST_NOALLOC_MAX_RAYS

EXPORTS = ["ptb_last_error", "ptb_version", "ptb_scene_create", "ptb_scene_destroy", "ptb_scene_info",
           "ptb_trace_primary", "ptb_trace_closest", "ptb_trace_paths_workspace", "ptb_trace_paths",
           "ptb_tonemap", "ptb_scatter_accum", "ptb_counters_read", "ptb_counters_reset", "ptb_pixel_count",
           "ptb_profile_enable", "ptb_profile_read", "ptb_fp32_peak", "ptb_trace_primary_workspace",
           "ptb_obj_parse", "ptb_obj_free", "ptb_obj_write", "ptb_bvh_export_bytes", "ptb_bvh_export", "ptb_bvh_import", "ptb_tonemap_scene", "ptb_pool_trim",
           "ptb_gather_peak", "ptb_gather_peak_smem", "ptb_kat", "ptb_scene_plane_culling",
           "ptb_trace_closest_noalloc", "ptb_trace_paths_at"]


class PtbError(RuntimeError):
    pass


class SceneDesc(C.Structure):
    _fields_ = [("n_tri", C.c_int32), ("n_tri_obj", C.c_int32), ("n_obj", C.c_int32), ("n_light_tri", C.c_int32),
                ("verts", C.c_void_p), ("normals", C.c_void_p), ("tri_obj", C.c_void_p), ("mat", C.c_void_p),
                ("light_areas", C.c_void_p), ("eye", C.c_double * 3), ("ortho", C.c_double * 4),
                ("ambient", C.c_double), ("light_color", C.c_double * 3), ("accel", C.c_int32), ("reserved", C.c_int32)]


class Pixels(C.Structure):
    _fields_ = [("width", C.c_int32), ("height", C.c_int32), ("shard", C.c_int32), ("n_shards", C.c_int32),
                ("tile", C.c_int32), ("n_pixels", C.c_int32), ("d_pixels", C.c_void_p)]


class RenderParams(C.Structure):
    _fields_ = [("px", Pixels), ("sample_begin", C.c_int32), ("sample_end", C.c_int32), ("bounces", C.c_int32),
                ("seed", C.c_uint32), ("precision", C.c_int32), ("flags", C.c_int32),
                ("samples_per_batch", C.c_int32), ("reserved", C.c_int32)]


class PathRecords(C.Structure):
    _fields_ = [("d_tri", C.c_void_p), ("d_contrib", C.c_void_p), ("d_k_after", C.c_void_p), ("d_lobe", C.c_void_p),
                ("d_occ", C.c_void_p), ("d_leak", C.c_void_p), ("d_contrib64", C.c_void_p), ("d_k_after64", C.c_void_p)]


class Counters(C.Structure):
    _fields_ = [(k, C.c_uint64) for k in ("closest_rays", "shadow_rays", "shaded", "samples", "retraced", "launches",
                                          "nodes_visited", "tris_tested", "nodes_visited_shadow", "tris_tested_shadow")]


PROF_KINDS = ["camera", "closest_hit", "fp64_retrace", "shade", "shadow", "reduce", "tonemap", "other", "bvh_build", "ray_sort", "bounce_fused"]

_lib = None


def lib():
    """Load libptb.so (built in-tree by __graft_entry__.build()).  Raises if it is not there."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise PtbError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'`; "
                           "there is no CPU fallback")
        L = C.CDLL(LIB_PATH)
        L.ptb_last_error.restype = C.c_char_p
        L.ptb_trace_paths_workspace.restype = C.c_size_t
        L.ptb_trace_primary_workspace.restype = C.c_size_t
        L.ptb_trace_primary_workspace.argtypes = [C.c_int64, C.c_int32]
        L.ptb_pixel_count.restype = C.c_int64
        L.ptb_scene_create.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.POINTER(C.c_void_p)]
        L.ptb_scene_destroy.argtypes = [C.c_void_p]
        L.ptb_bvh_export_bytes.restype = C.c_size_t
        L.ptb_bvh_export_bytes.argtypes = [C.c_void_p]
        L.ptb_bvh_export.argtypes = [C.c_void_p, C.c_void_p, C.c_size_t, C.c_void_p]
        L.ptb_bvh_import.argtypes = [C.POINTER(SceneDesc), C.c_void_p, C.c_size_t, C.c_void_p, C.POINTER(C.c_void_p)]
        L.ptb_scene_info.argtypes = [C.c_void_p, C.POINTER(C.c_int32), C.POINTER(C.c_int32), C.POINTER(C.c_double)]
        L.ptb_trace_primary.argtypes = [C.c_void_p, C.POINTER(Pixels), C.c_int32, C.c_int32, C.c_void_p, C.c_void_p,
                                        C.c_void_p, C.c_size_t, C.c_void_p]
        L.ptb_trace_closest.argtypes = [C.c_void_p, C.c_void_p, C.c_int64, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptb_trace_closest_noalloc.argtypes = L.ptb_trace_closest.argtypes
        L.ptb_trace_paths_workspace.argtypes = [C.POINTER(RenderParams), C.c_int32]
        L.ptb_trace_paths.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_size_t, C.c_void_p,
                                      C.POINTER(PathRecords), C.c_void_p]
        L.ptb_trace_paths_at.argtypes = [C.c_void_p, C.POINTER(RenderParams), C.c_void_p, C.c_size_t, C.c_void_p,
                                         C.POINTER(PathRecords), C.c_void_p, C.c_void_p]
        L.ptb_tonemap.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptb_tonemap_scene.argtypes = [C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_double, C.POINTER(C.c_double),
                                        C.c_void_p, C.c_void_p]
        L.ptb_scatter_accum.argtypes = [C.POINTER(Pixels), C.c_void_p, C.c_void_p, C.c_void_p]
        L.ptb_counters_read.argtypes = [C.c_void_p, C.POINTER(Counters), C.c_void_p]
        L.ptb_counters_reset.argtypes = [C.c_void_p, C.c_void_p]
        L.ptb_pixel_count.argtypes = [C.POINTER(Pixels)]
        L.ptb_profile_enable.argtypes = [C.c_int]
        L.ptb_fp32_peak.argtypes = [C.POINTER(C.c_double), C.c_void_p]
        L.ptb_obj_parse.argtypes = [C.c_char_p, C.POINTER(C.c_void_p), C.POINTER(C.c_int64), C.POINTER(C.c_void_p),
                                    C.POINTER(C.c_int64), C.POINTER(C.c_void_p), C.c_int]
        L.ptb_obj_free.argtypes = [C.c_void_p]
        L.ptb_obj_free.restype = None
        L.ptb_obj_write.argtypes = [C.c_char_p, C.c_void_p, C.c_int64, C.c_void_p, C.c_int64, C.c_char_p]
        L.ptb_profile_read.argtypes = [C.POINTER(C.c_double), C.POINTER(C.c_uint64)]
        L.ptb_gather_peak.argtypes = [C.c_size_t, C.POINTER(C.c_double), C.c_void_p]
        L.ptb_gather_peak_smem.argtypes = [C.c_size_t, C.POINTER(C.c_double), C.c_void_p]
        L.ptb_kat.argtypes = [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_void_p, C.c_int32, C.c_int32,
                              C.c_uint32, C.c_void_p]
        L.ptb_scene_plane_culling.argtypes = [C.c_void_p]
        _lib = L
    return _lib


def check(rc, what):
    if rc != 0:
        raise PtbError(f"{what} failed ({rc}): {lib().ptb_last_error().decode()}")


ERR_IO, ERR_PARSE, ERR_PARSE_INDEX, ERR_PARSE_IRREGULAR = -5, -6, -7, -8


def obj_parse(path, threads=0):
    """ptb_obj_parse -> (vertexes float64 [n,3], faces int64 [m,3], [cleaned text of skipped lines]).
    Raises what the reference's parser raises: OSError, ValueError ('i/j/k', bad numbers), IndexError."""
    import numpy as np
    L = lib()
    v, f, sk = C.c_void_p(), C.c_void_p(), C.c_void_p()
    nv, nf = C.c_int64(), C.c_int64()
    rc = L.ptb_obj_parse(os.fsencode(path), C.byref(v), C.byref(nv), C.byref(f), C.byref(nf), C.byref(sk), threads)
    if rc != 0:
        msg = L.ptb_last_error().decode()
        if rc == ERR_IO:
            raise FileNotFoundError(msg)
        if rc == ERR_PARSE:
            raise ValueError(msg)
        if rc == ERR_PARSE_INDEX:
            raise IndexError("string index out of range")
        if rc == ERR_PARSE_IRREGULAR:
            raise IndexError(msg)           # scene.Obj re-reads such files line by line, with the reference's semantics
        raise PtbError(f"ptb_obj_parse failed ({rc}): {msg}")
    try:
        verts = np.ctypeslib.as_array(C.cast(v, C.POINTER(C.c_double)), (nv.value * 3,)).reshape(-1, 3).copy() \
            if nv.value else np.zeros((0, 3), np.float64)
        faces = np.ctypeslib.as_array(C.cast(f, C.POINTER(C.c_int64)), (nf.value * 3,)).reshape(-1, 3).copy() \
            if nf.value else np.zeros((0, 3), np.int64)
        skipped = C.string_at(sk).decode(errors="replace").split("\n")[:-1] if sk.value else []
    finally:
        for p in (v, f, sk):
            if p.value:
                L.ptb_obj_free(p)
    return verts, faces, skipped


def obj_write(path, vertexes, faces, header=""):
    """ptb_obj_write: float64 [n,3] vertices and 0-based int64 [m,3] faces -> .obj text with repr-exact decimals."""
    import numpy as np
    v = np.ascontiguousarray(vertexes, np.float64).reshape(-1, 3)
    f = np.ascontiguousarray(faces, np.int64).reshape(-1, 3)
    rc = lib().ptb_obj_write(os.fsencode(path), v.ctypes.data, v.shape[0], f.ctypes.data, f.shape[0], header.encode())
    if rc == ERR_IO:
        raise OSError(lib().ptb_last_error().decode())
    check(rc, "ptb_obj_write")


def profile_enable(on=True):
    check(lib().ptb_profile_enable(1 if on else 0), "ptb_profile_enable")


def profile_read():
    """-> {kind: (milliseconds, launches)} accumulated since the last read (synchronises the device)."""
    ms = (C.c_double * len(PROF_KINDS))()
    n = (C.c_uint64 * len(PROF_KINDS))()
    check(lib().ptb_profile_read(ms, n), "ptb_profile_read")
    return {k: (float(ms[i]), int(n[i])) for i, k in enumerate(PROF_KINDS)}


def gather_peak_gfetch(buffer_bytes):
    """ptb_gather_peak: dependent random 64 B fetches per second (in 1e9) through a buffer of that size."""
    v = C.c_double()
    check(lib().ptb_gather_peak(int(buffer_bytes), C.byref(v), None), "ptb_gather_peak")
    return float(v.value)


def gather_peak_smem_gfetch(bytes_per_block=16 << 10):
    """ptb_gather_peak_smem: the same chase through records staged in shared memory (per block)."""
    v = C.c_double()
    check(lib().ptb_gather_peak_smem(int(bytes_per_block), C.byref(v), None), "ptb_gather_peak_smem")
    return float(v.value)


def fp32_peak_tflops():
    v = C.c_double()
    check(lib().ptb_fp32_peak(C.byref(v), None), "ptb_fp32_peak")
    return float(v.value)
