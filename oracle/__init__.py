"""oracle -- CPU checker for the WKV6 hot path.  TEST INFRASTRUCTURE, NOT PRODUCT CODE.

Only ``tests/``, ``__graft_entry__.smoke()`` and ``bench.py``'s ``cpu_baseline`` /
``--impl reference`` legs may import this package.  The product package
(``paper_accurate_fast_cheap_b200``) never does, and fails loudly without its CUDA library.

ctypes binding over ``libwkv6_oracle.so`` (oracle/wkv6_oracle.c), a C restatement of
/root/reference/wenet/rwkv_v6/cuda/wkv6_cuda.cu and wkv6state_cuda.cu.  ``*_f32`` keeps
the GPU kernel's operation order; ``*_f64`` is the truth used for tolerances.
"""
from __future__ import annotations

import ctypes
import os
import subprocess
from typing import Optional, Tuple

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
_SO = os.path.join(_HERE, "libwkv6_oracle.so")
_SRC = [os.path.join(_HERE, "wkv6_oracle.c"), os.path.join(_HERE, "wkv6_oracle_impl.h")]
_lib = None


def build(force: bool = False) -> str:
    """Compile the C restatement (gcc, OpenMP when available).  Returns the .so path."""
    if not force and os.path.exists(_SO) and all(
            os.path.getmtime(_SO) >= os.path.getmtime(s) for s in _SRC):
        return _SO
    base = ["-O3", "-march=x86-64-v3", "-fPIC", "-ffp-contract=off", "-fno-math-errno",
            "-shared", "-o", _SO, _SRC[0], "-lm"]
    last = None
    for cc in ("/usr/bin/gcc", os.environ.get("CC", "gcc"), "gcc", "cc"):
        for omp in (["-fopenmp"], []):
            try:
                subprocess.run([cc] + omp + base, check=True, capture_output=True, cwd=_HERE)
                return _SO
            except (subprocess.CalledProcessError, FileNotFoundError) as e:  # try next
                last = e
    raise RuntimeError(f"could not build the oracle: {last}")


def lib():
    global _lib
    if _lib is None:
        build()
        _lib = ctypes.CDLL(_SO)
        _lib.wkv6_oracle_num_threads.restype = ctypes.c_int
    return _lib


def num_threads() -> int:
    return int(lib().wkv6_oracle_num_threads())


def set_threads(n: int) -> None:
    lib().wkv6_oracle_set_threads(int(n))


def _real(prec: str):
    return {"f32": torch.float32, "f64": torch.float64}[prec]


def _p(t: Optional[torch.Tensor]):
    return ctypes.c_void_p(0 if t is None else t.data_ptr())


def _prep(t: Optional[torch.Tensor], dt) -> Optional[torch.Tensor]:
    if t is None:
        return None
    return t.detach().to(device="cpu", dtype=dt).contiguous()


def fwd(r, k, v, w, u, s0=None, rev: bool = False, want_state: bool = False,
        prec: str = "f32") -> Tuple[torch.Tensor, Optional[torch.Tensor]]:
    """y (and S_T) of the WKV6 scan.  Inputs [B,T,C], u [H,N], s0 [B,H,N,N] or None.
    Inputs of any dtype are upcast exactly; the result is in `prec`.
    Follows wkv6_cuda.cu:9-63 / wkv6state_cuda.cu:7-65."""
    dt = _real(prec)
    B, T, C = r.shape
    H = u.shape[0]
    r, k, v, w, u, s0 = (_prep(x, dt) for x in (r, k, v, w, u, s0))
    y = torch.empty(B, T, C, dtype=dt)
    sT = torch.empty(B, H, C // H, C // H, dtype=dt) if want_state else None
    getattr(lib(), f"wkv6_oracle_fwd_{prec}")(
        B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u), _p(s0), _p(y), _p(sT), int(rev))
    return y, sT


def bwd_ref(r, k, v, w, u, gy, prec: str = "f32"):
    """Literal restatement of kernel_backward_101/102/103/201 (wkv6_cuda.cu:66-263).
    Returns gr, gk, gv, gw [B,T,C] and gu [B,C] (per-batch partials)."""
    dt = _real(prec)
    B, T, C = r.shape
    H = u.shape[0]
    r, k, v, w, u, gy = (_prep(x, dt) for x in (r, k, v, w, u, gy))
    outs = [torch.empty(B, T, C, dtype=dt) for _ in range(4)] + [torch.empty(B, C, dtype=dt)]
    getattr(lib(), f"wkv6_oracle_bwd_ref_{prec}")(
        B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u), _p(gy), *[_p(o) for o in outs])
    return tuple(outs)


def bwd_state_ref(r, k, v, w, u, s0, gy, prec: str = "f32"):
    """Literal restatement of kernel_backward_111/222 (wkv6state_cuda.cu:68-293).
    Returns gr, gk, gv, gw, gu [B,C], gs [B,H,N,N]."""
    dt = _real(prec)
    B, T, C = r.shape
    H = u.shape[0]
    r, k, v, w, u, s0, gy = (_prep(x, dt) for x in (r, k, v, w, u, s0, gy))
    outs = [torch.empty(B, T, C, dtype=dt) for _ in range(4)] + [
        torch.empty(B, C, dtype=dt), torch.empty(B, H, C // H, C // H, dtype=dt)]
    getattr(lib(), f"wkv6_oracle_bwd_state_ref_{prec}")(
        B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u), _p(s0), _p(gy), *[_p(o) for o in outs])
    return tuple(outs)


def bwd(r, k, v, w, u, gy, s0=None, gsT=None, rev: bool = False, prec: str = "f64"):
    """General backward (extension: initial state, final-state gradient, reversed time).
    Returns gr, gk, gv, gw [B,T,C], gu [B,C], gs0 [B,H,N,N]."""
    dt = _real(prec)
    B, T, C = r.shape
    H = u.shape[0]
    r, k, v, w, u, gy, s0, gsT = (_prep(x, dt) for x in (r, k, v, w, u, gy, s0, gsT))
    outs = [torch.empty(B, T, C, dtype=dt) for _ in range(4)] + [
        torch.empty(B, C, dtype=dt), torch.empty(B, H, C // H, C // H, dtype=dt)]
    getattr(lib(), f"wkv6_oracle_bwd_{prec}")(
        B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u), _p(s0), _p(gy), _p(gsT),
        *[_p(o) for o in outs], int(rev))
    return tuple(outs)


def rel_err(a: torch.Tensor, b: torch.Tensor) -> float:
    """max|a-b| / max|b|  -- the tolerance definition of SURVEY.md section 8(c)."""
    a = a.detach().double().cpu()
    b = b.detach().double().cpu()
    den = float(b.abs().max())
    num = float((a - b).abs().max())
    if den == 0.0:
        return num
    return num / den
