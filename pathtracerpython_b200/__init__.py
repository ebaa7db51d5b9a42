"""B200-native path-tracing hot path of thiagoald/pathtracerpython: `.sdl` scene in, tone-mapped image
out, computed by hand-written sm_100a CUDA kernels behind a C ABI (include/ptb.h)."""
from ._lib import ACCEL_AUTO, ACCEL_BRUTE, ACCEL_LBVH, FLAG_COUNT_NODES, FLAG_NO_FP64_RETRACE, FLAG_SEPARATE_KERNELS, FLAG_PHYSICAL, FLAG_NO_ROULETTE, PREC_FP32, PREC_FP64, PtbError
from .scene import CORNELL_SDL, Obj, Scene
from .render import ProgressiveFrame, Renderer, release_workspace, render

__all__ = ["Scene", "Obj", "Renderer", "ProgressiveFrame", "render", "PtbError", "CORNELL_SDL", "PREC_FP32", "PREC_FP64", "ACCEL_AUTO",
           "release_workspace", "ACCEL_BRUTE", "ACCEL_LBVH", "FLAG_NO_FP64_RETRACE", "FLAG_SEPARATE_KERNELS", "FLAG_COUNT_NODES", "FLAG_PHYSICAL", "FLAG_NO_ROULETTE"]
