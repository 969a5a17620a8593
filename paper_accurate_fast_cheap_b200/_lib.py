"""ctypes binding of libwkv6_b200.so (include/wkv6_b200.h).  Fails loudly: there is no CPU fallback."""
from __future__ import annotations

import ctypes
import os

_HERE = os.path.dirname(os.path.abspath(__file__))
# WKV6_B200_LIB: another build of the same sources (the phase-trace build of build.py --trace, A/B variants of a kernel)
LIB_PATH = os.environ.get("WKV6_B200_LIB") or os.path.join(_HERE, "libwkv6_b200.so")
TORCH_LIB_PATH = os.path.join(_HERE, "libwkv6_b200_torch.so")

DTYPE_BF16, DTYPE_F32 = 0, 1
BI_L2R, BI_R2L = 1, 2
ALGO_AUTO, ALGO_SCAN, ALGO_CHUNK = 0, 1, 2

# every symbol include/wkv6_b200.h declares
SYMBOLS = (
    "wkv6_b200_abi_version", "wkv6_b200_last_error", "wkv6_b200_launch_count", "wkv6_b200_set_algo",
    "wkv6_b200_get_algo", "wkv6_b200_debug_trace", "wkv6_b200_forward_scratch_bytes", "wkv6_b200_debug_trace_ctas", "wkv6_b200_checkpoint_bytes", "wkv6_b200_forward", "wkv6_b200_forward_static_scratch", "wkv6_b200_backward_workspace_bytes", "wkv6_b200_backward_scratch_bytes", "wkv6_b200_backward",
    "wkv6_b200_bi_forward", "wkv6_b200_bi_backward", "wkv6_b200_combine", "wkv6_b200_tmix_mix", "wkv6_b200_tmix_mix_backward", "wkv6_b200_ln_pair_cat", "wkv6_b200_convmod_core",
)

_lib = None
_vp, _i, _sz = ctypes.c_void_p, ctypes.c_int, ctypes.c_size_t


class Wkv6Error(RuntimeError):
    pass


def lib() -> ctypes.CDLL:
    """Loads the CUDA library; raises if it has not been built (python -m paper_accurate_fast_cheap_b200.build)."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise Wkv6Error(f"{LIB_PATH} is missing: build it with `python -m paper_accurate_fast_cheap_b200.build` "
                        "(there is no CPU or PyTorch fallback for the WKV6 path)")
    L = ctypes.CDLL(LIB_PATH, mode=ctypes.RTLD_GLOBAL)
    L.wkv6_b200_abi_version.restype = _i
    L.wkv6_b200_last_error.restype = ctypes.c_char_p
    L.wkv6_b200_launch_count.restype = ctypes.c_longlong
    L.wkv6_b200_set_algo.argtypes = [_i]
    L.wkv6_b200_get_algo.restype = _i
    L.wkv6_b200_debug_trace.argtypes = [_i, _vp]
    L.wkv6_b200_debug_trace.restype = _i
    L.wkv6_b200_forward_scratch_bytes.argtypes = [_i] * 6
    L.wkv6_b200_forward_scratch_bytes.restype = _sz
    L.wkv6_b200_debug_trace_ctas.argtypes = [_vp]
    L.wkv6_b200_debug_trace_ctas.restype = _i
    L.wkv6_b200_checkpoint_bytes.argtypes = [_i] * 4
    L.wkv6_b200_checkpoint_bytes.restype = _sz
    L.wkv6_b200_forward.argtypes = [_i] * 5 + [_vp] * 9 + [_i, _vp, _sz, _vp]
    L.wkv6_b200_forward.restype = _i
    L.wkv6_b200_forward_static_scratch.argtypes = [_i] * 5 + [_vp] * 9 + [_i, _vp]
    L.wkv6_b200_forward_static_scratch.restype = _i
    L.wkv6_b200_backward_workspace_bytes.argtypes = [_i] * 6
    L.wkv6_b200_backward_workspace_bytes.restype = _sz
    L.wkv6_b200_backward_scratch_bytes.argtypes = [_i] * 6
    L.wkv6_b200_backward_scratch_bytes.restype = _sz
    L.wkv6_b200_backward.argtypes = [_i] * 5 + [_vp] * 15 + [_i, _vp, _sz, _vp]
    L.wkv6_b200_backward.restype = _i
    L.wkv6_b200_bi_forward.argtypes = [_i] * 5 + [_vp] * 18 + [_i, _vp, _sz, _vp]
    L.wkv6_b200_bi_forward.restype = _i
    L.wkv6_b200_bi_backward.argtypes = [_i] * 5 + [_vp] * 30 + [_i, _vp, _sz, _vp]
    L.wkv6_b200_bi_backward.restype = _i
    L.wkv6_b200_combine.argtypes = [_i, _sz, _vp, _vp, ctypes.c_float, _i, _vp, _vp]
    L.wkv6_b200_combine.restype = _i
    L.wkv6_b200_tmix_mix.argtypes = [_i] * 4 + [_vp] * 8 + [_i] + [_vp] * 6
    L.wkv6_b200_tmix_mix.restype = _i
    L.wkv6_b200_tmix_mix_backward.argtypes = [_i] * 4 + [_vp] * 10 + [_i] + [_vp] * 4
    L.wkv6_b200_tmix_mix_backward.restype = _i
    L.wkv6_b200_ln_pair_cat.argtypes = [_i, ctypes.c_longlong, _i] + [_vp] * 6 + [ctypes.c_float, _vp, _vp]
    L.wkv6_b200_ln_pair_cat.restype = _i
    L.wkv6_b200_convmod_core.argtypes = [_i] * 7 + [_vp] * 5 + [ctypes.c_float, _vp, _vp]
    L.wkv6_b200_convmod_core.restype = _i
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise Wkv6Error(f"wkv6_b200 error {rc}: {lib().wkv6_b200_last_error().decode()}")


def launch_count() -> int:
    return int(lib().wkv6_b200_launch_count())


def set_algo(algo: int) -> None:
    lib().wkv6_b200_set_algo(int(algo))


_torch_ops_loaded = False


def load_torch_ops() -> None:
    """Registers torch.ops.wkv6.{forward,backward}[_fp32|_fp64] (the reference's dispatcher namespace)."""
    global _torch_ops_loaded
    if _torch_ops_loaded:
        return
    import torch
    lib()
    if not os.path.exists(TORCH_LIB_PATH):
        raise Wkv6Error(f"{TORCH_LIB_PATH} is missing: build it with `python -m paper_accurate_fast_cheap_b200.build`")
    torch.ops.load_library(TORCH_LIB_PATH)
    _torch_ops_loaded = True
