"""rust_raytracer_b200 — B200-native path-tracing backend behind the rust-raytracer scene API.

Host side: a mirror of the reference's construction API (api.py) that flattens scenes into the tables of
include/b200rt.h; device side: hand-written sm_100a CUDA kernels behind the C ABI (csrc/).  Importing the package
does not load the library; the first call does, and fails loudly if libb200rt.so has not been built.
"""
from . import _ffi as ffi  # noqa: F401
from .api import *  # noqa: F401,F403
from .engine import Context, Scene, make_params, copy_params, render_multi, render_sppm_multi, compile_stats, bvh_quality, HIT_DTYPE  # noqa: F401
