"""autograd.Function layer: same names and call signatures as the reference (model.py:108-214) plus the
state / reversed / bidirectional superset."""
from __future__ import annotations

from typing import Optional

import torch

from . import _lib, ops

HEAD_SIZE = ops.HEAD_SIZE


class _WKV6Base(torch.autograd.Function):
    """r, k, v, w [B,T,C], u [H,N] -> y [B,T,C].  S_0 = 0, no state out: the reference's live op."""
    DTYPE = torch.bfloat16

    @classmethod
    def _fwd(cls, ctx, r, k, v, w, u):
        with torch.no_grad():
            B, T, C = r.size()
            H = C // HEAD_SIZE
            assert C % HEAD_SIZE == 0
            for t in (r, k, v, w, u):
                assert t.dtype == cls.DTYPE
                assert t.is_contiguous()
            ctx.B, ctx.T, ctx.C, ctx.H = B, T, C, H
            need = any(ctx.needs_input_grad)
            ck = ops.checkpoint_buffer(r) if need else None
            ctx.has_ck = ck is not None
            ctx.save_for_backward(r, k, v, w, u, *((ck,) if ck is not None else ()))
            y, _ = ops.wkv6_forward_raw(r, k, v, w, u, ckpt=ck)
            return y

    @classmethod
    def _bwd(cls, ctx, gy):
        with torch.no_grad():
            assert gy.dtype == cls.DTYPE
            gy = gy.contiguous()
            r, k, v, w, u = ctx.saved_tensors[:5]
            ck = ctx.saved_tensors[5] if ctx.has_ck else None
            gr, gk, gv, gw, gu, _ = ops.wkv6_backward_raw(r, k, v, w, u, gy, ckpt=ck)
            gu = torch.sum(gu, 0).view(ctx.H, ctx.C // ctx.H)      # model.py:151
            return gr, gk, gv, gw, gu


class WKV_6(_WKV6Base):
    DTYPE = torch.bfloat16

    @staticmethod
    def forward(ctx, r, k, v, w, u):
        return WKV_6._fwd(ctx, r, k, v, w, u)

    @staticmethod
    def backward(ctx, gy):
        return WKV_6._bwd(ctx, gy)


class WKV_6_FP32(_WKV6Base):
    DTYPE = torch.float32

    @staticmethod
    def forward(ctx, r, k, v, w, u):
        return WKV_6_FP32._fwd(ctx, r, k, v, w, u)

    @staticmethod
    def backward(ctx, gy):
        return WKV_6_FP32._bwd(ctx, gy)


def RUN_CUDA_RWKV6(r, k, v, w, u):
    return WKV_6.apply(r, k, v, w, u)


def RUN_CUDA_RWKV6_FP32(r, k, v, w, u):
    return WKV_6_FP32.apply(r, k, v, w, u)


class WKV6State(torch.autograd.Function):
    """(r,k,v,w,u,s0) -> (y, S_T); s0 may be None; `reverse` scans t = T-1..0 in place.
    Superset of WKV_6STATE (model.py:54-103): fp32 state, final state out, gradient through both."""

    @staticmethod
    def forward(ctx, r, k, v, w, u, s0, reverse):
        with torch.no_grad():
            ctx.reverse = bool(reverse)
            ctx.has_s0 = s0 is not None
            ck = ops.checkpoint_buffer(r) if any(ctx.needs_input_grad) else None
            ctx.has_ck = ck is not None
            y, sT = ops.wkv6_forward_raw(r, k, v, w, u, s0=s0, reverse=reverse, want_state=True, ckpt=ck)
            ctx.save_for_backward(r, k, v, w, u, *((s0,) if s0 is not None else ()), *((ck,) if ck is not None else ()))
            ctx.H = u.shape[0]
            return y, sT

    @staticmethod
    def backward(ctx, gy, gsT):
        with torch.no_grad():
            saved = ctx.saved_tensors
            r, k, v, w, u = saved[:5]
            s0 = saved[5] if ctx.has_s0 else None
            ck = saved[-1] if ctx.has_ck else None
            if gy is None:
                gy = torch.zeros_like(r)
            gsT = None if gsT is None else gsT.contiguous().float()
            gr, gk, gv, gw, gu, gs0 = ops.wkv6_backward_raw(r, k, v, w, u, gy.contiguous(), s0=s0, gsT=gsT,
                                                            reverse=ctx.reverse, want_gs0=ctx.has_s0, ckpt=ck)
            gu = torch.sum(gu, 0).view(ctx.H, -1)
            return gr, gk, gv, gw, gu, gs0, None


def wkv6(r, k, v, w, u, s0: Optional[torch.Tensor] = None, reverse: bool = False):
    """y, S_T = WKV6(r,k,v,w,u | S_0), differentiable in everything."""
    return WKV6State.apply(r, k, v, w, u, s0, reverse)


class BiWKV6(torch.autograd.Function):
    """Both directions in one launch per pass.  Inputs: the L2R set (r,k,v,w,u) and the R2L set; `flags`
    (BI_L2R | BI_R2L) implements direction dropout: a dropped direction is not read and returns None / no grad."""

    @staticmethod
    def forward(ctx, flags, r_f, k_f, v_f, w_f, u_f, r_b, k_b, v_b, w_b, u_b):
        with torch.no_grad():
            ctx.flags = int(flags)
            f, b = (r_f, k_f, v_f, w_f, u_f), (r_b, k_b, v_b, w_b, u_b)
            run0 = f if (ctx.flags & _lib.BI_L2R) else b
            ck = ops.checkpoint_buffer(run0[0], 2) if any(ctx.needs_input_grad) else None
            ctx.has_ck = ck is not None
            y_f, y_b, _, _ = ops.bi_forward_raw(f, b, flags=ctx.flags, ckpt=ck)
            ctx.save_for_backward(*f, *b, *((ck,) if ck is not None else ()))
            run_f, run_b = bool(ctx.flags & _lib.BI_L2R), bool(ctx.flags & _lib.BI_R2L)
            ref = r_f if run_f else r_b
            # autograd wants tensors: a dropped direction yields an empty placeholder, marked non-differentiable
            if y_f is None:
                y_f = ref.new_empty(0)
                ctx.mark_non_differentiable(y_f)
            if y_b is None:
                y_b = ref.new_empty(0)
                ctx.mark_non_differentiable(y_b)
            return y_f, y_b

    @staticmethod
    def backward(ctx, gy_f, gy_b):
        with torch.no_grad():
            t = ctx.saved_tensors
            f, b = t[:5], t[5:10]
            ck = t[10] if ctx.has_ck else None
            flags = ctx.flags
            run_f, run_b = bool(flags & _lib.BI_L2R), bool(flags & _lib.BI_R2L)
            if run_f:
                gy_f = (torch.zeros_like(f[0]) if gy_f is None else gy_f).contiguous()
            if run_b:
                gy_b = (torch.zeros_like(b[0]) if gy_b is None else gy_b).contiguous()
            of, ob = ops.bi_backward_raw(f, b, gy_f if run_f else None, gy_b if run_b else None, flags=flags, ckpt=ck)
            H = (f[4] if run_f else b[4]).shape[0]

            def pack(o, run):
                if not run:
                    return (None,) * 5
                return o[0], o[1], o[2], o[3], torch.sum(o[4], 0).view(H, -1)

            return (None,) + pack(of, run_f) + pack(ob, run_b)


def bi_wkv6(fwd_set, bwd_set, run_l2r: bool = True, run_r2l: bool = True):
    """y_f, y_b of the two directions (None where dropped).  Each set is (r,k,v,w,u)."""
    flags = (_lib.BI_L2R if run_l2r else 0) | (_lib.BI_R2L if run_r2l else 0)
    assert flags, "bi_wkv6: at least one direction must run"
    y_f, y_b = BiWKV6.apply(flags, *fwd_set, *bwd_set)
    return (y_f if run_l2r else None), (y_b if run_r2l else None)


class Combine(torch.autograd.Function):
    """out = (a + b) * scale as float32 in one fused pass (rwkv_wrapper_bidirectional.py:48-56).
    `round_sum` rounds a + b to bf16 first, like the reference's bf16 add."""

    @staticmethod
    def forward(ctx, a, b, scale, round_sum):
        ctx.scale, ctx.dt = float(scale), a.dtype
        with torch.no_grad():
            return ops.combine(a.contiguous(), b.contiguous(), ctx.scale, round_sum)

    @staticmethod
    def backward(ctx, g):
        ga = (g * ctx.scale).to(ctx.dt)
        return ga, ga, None, None


def combine(a, b, scale: float = 0.5, round_sum: bool = False, out_float: bool = True):
    out = Combine.apply(a, b, scale, bool(round_sum) and a.dtype == torch.bfloat16)
    return out if out_float else out.to(a.dtype)
