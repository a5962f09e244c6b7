"""ctypes binding of libb200ad.so (the C ABI declared in include/b200ad.h).

The product path has no CPU fallback: if the shared library is missing, or a kernel call fails, this module
raises.  Build the library with ``python -c "import __graft_entry__ as g; g.build()"`` (or ``make -C csrc``).
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from ctypes import c_char_p, c_float, c_int, c_longlong, c_void_p

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("B200AD_LIB") or os.path.join(_HERE, "libb200ad.so")   # B200AD_LIB: profiling build
CSRC_DIR = os.path.join(_HERE, "csrc")

_lib = None

# name -> argtypes (restype is int unless listed in _RESTYPES); mirrors include/b200ad.h one to one
_P, _I, _L, _F = c_void_p, c_int, c_longlong, c_float
SIGNATURES = {
    "b200ad_last_error": [],
    "b200ad_version": [],
    "b200ad_launch_count": [],
    "b200ad_debug_buffer": [_P],
    "b200ad_gemm": [_I, _P, _L, _P, _I, _I, _I, _P, _P, _I, _I, _P, _L, _P, _L, _P, _L, _I, _P],
    "b200ad_conv3x3": [_I, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _I, _I, _P, _L, _P, _L, _P, _L, _P],
    "b200ad_gn_nblk": [_I],
    "b200ad_stats_slots": [_I, _I, _I],
    "b200ad_gn_stats_ws_bytes": [_I, _I, _I],
    "b200ad_gemm_stats": [_I, _P, _L, _P, _I, _I, _I, _P, _P, _I, _I, _P, _L, _P, _L, _P, _I, _P],
    "b200ad_conv3x3_stats": [_I, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P, _I, _I, _P, _L, _P, _L, _P, _P],
    "b200ad_conv3x3_sc": [_I, _P, _P, _I, _I, _I, _I, _I, _P, _I, _P, _P, _L, _P, _L, _P, _P],
    "b200ad_upconv3x3": [_I, _P, _P, _I, _I, _I, _I, _I, _P, _P, _L, _P, _L, _P, _P],
    "b200ad_gn_apply_stats": [_I, _P, _I, _P, _I, _P, _I, _P, _I, _I, _I, _I, _I, _F, _P, _P, _I, _P, _P, _P, _P],
    "b200ad_gn_counters_nonzero": [],
    "b200ad_gn_stats": [_P, _I, _P, _I, _I, _I, _I, _P, _P],
    "b200ad_gn_apply": [_I, _P, _I, _P, _I, _I, _I, _I, _P, _I, _F, _P, _P, _I, _P, _P, _P],
    "b200ad_group_norm": [_I, _P, _I, _P, _I, _I, _I, _I, _P, _I, _F, _P, _P, _I, _P, _P, _P],
    "b200ad_layernorm": [_I, _P, _I, _I, _P, _P, _F, _P, _I, _I, _P, _P],
    "b200ad_cast": [_I, _P, _I, _I, _I, _I, _I, _P, _P],
    "b200ad_attention": [_I, _P, _L, _P, _L, _P, _L, _I, _I, _I, _I, _I, _I, _I, _P, _L, _P],
    "b200ad_attention2": [_I, _P, _L, _P, _L, _P, _L, _I, _I, _I, _I, _I, _I, _I, _I, _P, _L, _P],
    "b200ad_temporal_attention": [_I, _P, _I, _I, _I, _I, _I, _P, _P],
    "b200ad_timestep_sinusoid": [_P, _I, _I, _P, _P],
    "b200ad_small_linear": [_P, _I, _I, _P, _P, _I, _I, _I, _P, _P],
    "b200ad_conv_in": [_P, _I, _I, _I, _I, _I, _P, _P, _I, _P, _P],
    "b200ad_conv_out": [_I, _P, _I, _I, _I, _I, _I, _P, _P, _I, _P, _P],
    "b200ad_conv_in_affine": [_P, _I, _I, _I, _I, _I, _P, _P, _I, _P, _P],
    "b200ad_eltwise": [_I, _P, _P, _L, _L, _L, _P, _P],
    "b200ad_embed_tokens": [_P, _L, _I, _I, _P, _P, _I, _P, _P],
    "b200ad_attention_small": [_I, _P, _L, _P, _L, _P, _L, _I, _I, _I, _I, _I, _P, _L, _P],
    "b200ad_softmax_rows": [_I, _P, _L, _I, _L, _F, _P, _L, _P],
    "b200ad_video_to_u8": [_P, _I, _I, _I, _I, _P, _P],
    "b200ad_metric_blocks": [],
    "b200ad_video_pair_sums": [_P, _I, _L, _P, _P, _P],
    "b200ad_static_max_frames": [],
    "b200ad_video_pair_sums_static": [_P, _I, _L, _P, _P],
    "b200ad_video_warp_error": [_P, _I, _I, _I, _I, _P, _P, _P, _P],
    "b200ad_video_gray_u8": [_P, _L, _I, _P, _P],
    "b200ad_lpips_input": [_P, _L, _P, _P],
    "b200ad_conv2d_direct": [_P, _I, _I, _I, _I, _P, _P, _I, _I, _I, _I, _I, _I, _P, _P],
    "b200ad_maxpool2d": [_P, _I, _I, _I, _I, _I, _I, _P, _P],
    "b200ad_lpips_layer": [_P, _I, _I, _I, _P, _P, _P, _P],
    "b200ad_cfg_step": [_P, _P, _P, _F, _F, _F, _F, _L, _P, _P, _I, _P],
    "b200ad_cfg_step_dev": [_P, _P, _P, _P, _L, _P, _P, _I, _P],
}
_RESTYPES = {"b200ad_last_error": c_char_p, "b200ad_launch_count": c_longlong, "b200ad_gn_stats_ws_bytes": c_longlong}


class B200ADError(RuntimeError):
    pass


def build(verbose: bool = False) -> str:
    """Compile csrc/*.cu for sm_100a into libb200ad.so (nvcc cross-compiles without a GPU)."""
    r = subprocess.run(["make", "-C", CSRC_DIR, "-j8"], capture_output=True, text=True)
    if verbose or r.returncode != 0:
        print(r.stdout[-4000:])
        print(r.stderr[-4000:])
    if r.returncode != 0:
        raise B200ADError("building libb200ad.so failed")
    return LIB_PATH


def load():
    """Load the shared library (once) and attach argtypes."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise B200ADError(
            f"{LIB_PATH} not found: the CUDA extension is required (there is no CPU fallback). "
            "Run `python -c 'import __graft_entry__ as g; g.build()'` first.")
    lib = ctypes.CDLL(LIB_PATH)
    for name, argtypes in SIGNATURES.items():
        fn = getattr(lib, name)          # AttributeError if the library does not export a declared symbol
        fn.argtypes = argtypes
        fn.restype = _RESTYPES.get(name, c_int)
    _lib = lib
    return lib


def check(rc: int, what: str):
    if rc != 0:
        msg = load().b200ad_last_error()
        raise B200ADError(f"{what} failed (rc={rc}): {msg.decode() if msg else '?'}")


def launch_count() -> int:
    return int(load().b200ad_launch_count())
