"""Host side of the WKV6 operator: the reference's autograd.Function API on top of the C ABI.

Mirrors wenet/rwkv_v6/src/model.py:108-214 (WKV_6, WKV_6_FP32, RUN_CUDA_RWKV6[_FP32]): same names, argument
order, dtype/contiguity asserts and the same `gu.sum(0).view(H, N)` post-op.  Adds the superset the north star
asks for: initial state in, final state out, reversed (flip-free) direction, and both directions in one launch
with direction-dropout flags (`bi_wkv6`).  PyTorch is used for memory and streams only; all arithmetic happens in
libwkv6_b200.so.  There is no CPU path: CPU tensors raise.
"""
from __future__ import annotations

from typing import Optional, Tuple

import torch

from . import _lib

HEAD_SIZE = 64


def _dtype_code(t: torch.Tensor) -> int:
    if t.dtype == torch.bfloat16:
        return _lib.DTYPE_BF16
    if t.dtype == torch.float32:
        return _lib.DTYPE_F32
    raise TypeError(f"wkv6: unsupported dtype {t.dtype} (bfloat16 and float32 only)")


def _p(t: Optional[torch.Tensor]):
    return None if t is None else t.data_ptr()


def _stream(t: torch.Tensor):
    return torch.cuda.current_stream(t.device).cuda_stream


def _check_inputs(r, k, v, w, u, extra=()):
    assert r.is_cuda, "wkv6: CUDA tensors only (no CPU fallback in this build)"
    B, T, C = r.size()
    assert C % HEAD_SIZE == 0
    H = C // HEAD_SIZE
    for t in (r, k, v, w) + tuple(extra):
        assert t.dtype == r.dtype and t.device == r.device
        assert t.is_contiguous()
        assert t.size() == r.size()
    assert u.dtype == r.dtype and u.is_contiguous() and u.numel() == C
    return B, T, C, H


_segmentation = True


def set_segmentation(enabled: bool) -> None:
    """Test / A-B switch: with False no scratch is handed to the forward, so every launch runs one CTA per sequence."""
    global _segmentation
    _segmentation = bool(enabled)


_balanced = True


def set_balanced(enabled: bool) -> None:
    """Test / A-B switch: with False the chunk backward of more sequences than SMs gets no scratch and runs one CTA per
    sequence instead of the balanced persistent grid (hand-over records in the scratch)."""
    global _balanced
    _balanced = bool(enabled)


def _scratch(t: torch.Tensor, ndir: int):
    """Scratch for the time-segmented forward of few, long sequences (include/wkv6_b200.h): allocated per call from the
    caching allocator -- stream-ordered, so concurrent forwards on different streams never share a buffer -- and only for
    shapes the library would segment (wkv6_b200_forward_scratch_bytes returns 0 otherwise)."""
    B, T, C = t.size()
    if not _segmentation or T < 2048:
        return None, 0
    nb = _lib.lib().wkv6_b200_forward_scratch_bytes(_dtype_code(t), B, T, C, C // HEAD_SIZE, ndir)
    if nb == 0:
        return None, 0
    return torch.empty(nb, device=t.device, dtype=torch.uint8), nb


def _check_state(s, B, H, dev):
    if s is not None:
        assert s.dtype == torch.float32 and s.is_contiguous() and s.device == dev
        assert tuple(s.shape) == (B, H, HEAD_SIZE, HEAD_SIZE)


# --------------------------------------------------------------------------------------------------
# raw (no autograd) calls into the C ABI
# --------------------------------------------------------------------------------------------------
def checkpoint_buffer(r, ndir=1):
    """Buffer for the chunk-start state checkpoints of `ndir` directions (bf16 tcgen05 path only), or None."""
    if r.dtype != torch.bfloat16 or _lib.lib().wkv6_b200_get_algo() == _lib.ALGO_SCAN or r.size(1) == 0:
        return None
    B, T, C = r.size()
    nb = _lib.lib().wkv6_b200_checkpoint_bytes(B, T, C, C // HEAD_SIZE)
    return torch.empty((ndir, nb), device=r.device, dtype=torch.uint8)


def wkv6_forward_raw(r, k, v, w, u, s0=None, reverse=False, want_state=False, ckpt=None):
    B, T, C, H = _check_inputs(r, k, v, w, u)
    _check_state(s0, B, H, r.device)
    y = torch.empty_like(r)
    sT = torch.empty((B, H, HEAD_SIZE, HEAD_SIZE), device=r.device, dtype=torch.float32) if want_state else None
    with torch.cuda.device(r.device):
        scratch, nb = _scratch(r, 1)
        _lib.check(_lib.lib().wkv6_b200_forward(_dtype_code(r), B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u),
                                                _p(s0), _p(y), _p(sT), _p(ckpt), int(bool(reverse)), _p(scratch), nb,
                                                _stream(r)))
    return y, sT


def _bwd_workspace_bytes(L, r, B, T, C, H, ndir, ckpt) -> int:
    """Without checkpoints: everything the backward needs.  With checkpoints: only the scratch of the time-segmented sweep
    (few, long sequences; `set_segmentation(False)` turns it off) or of the balanced persistent grid (more sequences than
    SMs; `set_balanced(False)`); 0 for every other shape."""
    # at most half as many sequences as SMs: the time-segmented sweep; more sequences than SMs: the balanced persistent grid
    few_long = 2 * ndir * B * H <= torch.cuda.get_device_properties(r.device).multi_processor_count
    want_scratch = (_segmentation and T >= 2048) if few_long else _balanced
    if ckpt is None:
        nb = L.wkv6_b200_backward_workspace_bytes(_dtype_code(r), B, T, C, H, ndir)
        return nb if want_scratch else nb - L.wkv6_b200_backward_scratch_bytes(_dtype_code(r), B, T, C, H, ndir)
    if not want_scratch:
        return 0
    return L.wkv6_b200_backward_scratch_bytes(_dtype_code(r), B, T, C, H, ndir)


def wkv6_backward_raw(r, k, v, w, u, gy, s0=None, gsT=None, reverse=False, want_gs0=False, ckpt=None):
    """Returns gr, gk, gv, gw [B,T,C], gu [B,C] (per-batch partials, like the reference), gs0 or None."""
    B, T, C, H = _check_inputs(r, k, v, w, u, extra=(gy,))
    _check_state(s0, B, H, r.device)
    _check_state(gsT, B, H, r.device)
    gr, gk, gv, gw = (torch.empty_like(r) for _ in range(4))
    gu = torch.empty((B, C), device=r.device, dtype=r.dtype)
    gs0 = torch.empty((B, H, HEAD_SIZE, HEAD_SIZE), device=r.device, dtype=torch.float32) if want_gs0 else None
    L = _lib.lib()
    ws_bytes = _bwd_workspace_bytes(L, r, B, T, C, H, 1, ckpt)
    ws = torch.empty(ws_bytes, device=r.device, dtype=torch.uint8) if ws_bytes else None
    with torch.cuda.device(r.device):
        _lib.check(L.wkv6_b200_backward(_dtype_code(r), B, T, C, H, _p(r), _p(k), _p(v), _p(w), _p(u), _p(s0),
                                        _p(gy), _p(gsT), _p(ckpt), _p(gr), _p(gk), _p(gv), _p(gw), _p(gu), _p(gs0),
                                        int(bool(reverse)), _p(ws), ws_bytes, _stream(r)))
    return gr, gk, gv, gw, gu, gs0


def bi_forward_raw(f, b, flags=_lib.BI_L2R | _lib.BI_R2L, s0_f=None, s0_b=None, want_state=False, ckpt=None):
    """f, b: (r, k, v, w, u) of the L2R / R2L direction.  One launch.  Returns (y_f, y_b, sT_f, sT_b);
    outputs of a direction whose flag bit is clear are None."""
    ref = f if (flags & _lib.BI_L2R) else b
    B, T, C, H = _check_inputs(*ref)
    run_f, run_b = bool(flags & _lib.BI_L2R), bool(flags & _lib.BI_R2L)
    if run_f and run_b:
        _check_inputs(*b)
        assert b[0].size() == f[0].size() and b[0].dtype == f[0].dtype
    dev, dt = ref[0].device, ref[0].dtype
    mk = lambda: torch.empty((B, T, C), device=dev, dtype=dt)
    mks = lambda: torch.empty((B, H, HEAD_SIZE, HEAD_SIZE), device=dev, dtype=torch.float32)
    y_f = mk() if run_f else None
    y_b = mk() if run_b else None
    sT_f = mks() if (run_f and want_state) else None
    sT_b = mks() if (run_b and want_state) else None
    none5 = (None,) * 5
    fa = tuple(_p(t) for t in f) if run_f else none5
    ba = tuple(_p(t) for t in b) if run_b else none5
    ck_f = ckpt[0] if (ckpt is not None and run_f) else None
    ck_b = ckpt[1] if (ckpt is not None and run_b) else None
    with torch.cuda.device(dev):
        scratch, nb = _scratch(ref[0], int(run_f) + int(run_b))
        _lib.check(_lib.lib().wkv6_b200_bi_forward(_dtype_code(ref[0]), B, T, C, H,
                                                   *fa, _p(s0_f), _p(y_f), _p(sT_f), _p(ck_f),
                                                   *ba, _p(s0_b), _p(y_b), _p(sT_b), _p(ck_b), int(flags),
                                                   _p(scratch), nb, _stream(ref[0])))
    return y_f, y_b, sT_f, sT_b


def bi_backward_raw(f, b, gy_f, gy_b, flags=_lib.BI_L2R | _lib.BI_R2L, s0_f=None, s0_b=None, gsT_f=None, gsT_b=None,
                    want_gs0=False, ckpt=None):
    """Backward of bi_forward_raw: two launches (sweep A, sweep B) covering both directions.
    Returns two tuples (gr, gk, gv, gw, gu[B,C], gs0) -- None for a direction that did not run."""
    run_f, run_b = bool(flags & _lib.BI_L2R), bool(flags & _lib.BI_R2L)
    ref = f if run_f else b
    B, T, C, H = _check_inputs(*ref)
    dev, dt = ref[0].device, ref[0].dtype

    def outs(run):
        if not run:
            return (None,) * 6
        o = [torch.empty((B, T, C), device=dev, dtype=dt) for _ in range(4)]
        o.append(torch.empty((B, C), device=dev, dtype=dt))
        o.append(torch.empty((B, H, HEAD_SIZE, HEAD_SIZE), device=dev, dtype=torch.float32) if want_gs0 else None)
        return tuple(o)

    of, ob = outs(run_f), outs(run_b)
    L = _lib.lib()
    ws_bytes = _bwd_workspace_bytes(L, ref[0], B, T, C, H, int(run_f) + int(run_b), ckpt)
    ck_f = ckpt[0] if (ckpt is not None and run_f) else None
    ck_b = ckpt[1] if (ckpt is not None and run_b) else None
    ws = torch.empty(ws_bytes, device=dev, dtype=torch.uint8) if ws_bytes else None
    none5 = (None,) * 5
    fa = tuple(_p(t) for t in f) if run_f else none5
    ba = tuple(_p(t) for t in b) if run_b else none5
    with torch.cuda.device(dev):
        _lib.check(L.wkv6_b200_bi_backward(
            _dtype_code(ref[0]), B, T, C, H,
            *fa, _p(s0_f), _p(gy_f), _p(gsT_f), _p(ck_f), *[_p(t) for t in of],
            *ba, _p(s0_b), _p(gy_b), _p(gsT_b), _p(ck_b), *[_p(t) for t in ob],
            int(flags), _p(ws), ws_bytes, _stream(ref[0])))
    return of, ob


def combine(a: torch.Tensor, b: Optional[torch.Tensor], scale: float = 0.5, round_sum: bool = False) -> torch.Tensor:
    """(a + b) * scale as float32 in one pass (rwkv_wrapper_bidirectional.py:48-56)."""
    assert a.is_cuda and a.is_contiguous() and (b is None or (b.is_contiguous() and b.shape == a.shape and b.dtype == a.dtype))
    out = torch.empty(a.shape, device=a.device, dtype=torch.float32)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().wkv6_b200_combine(_dtype_code(a), a.numel(), _p(a), _p(b), float(scale), int(bool(round_sum)),
                                                _p(out), _stream(a)))
    return out


def ln_pair_cat(y_f: torch.Tensor, y_b: torch.Tensor, ln_f, ln_b) -> torch.Tensor:
    """[ln_f(y_f) || ln_b(y_b)] as one kernel: y_* [B,T,C] -> [B,T,2C] (include/wkv6_b200.h: wkv6_b200_ln_pair_cat).
    ln_f / ln_b: the two nn.LayerNorm modules (model.py:262, one per direction).  No autograd graph is recorded."""
    assert y_f.is_cuda and y_f.is_contiguous() and y_b.is_contiguous() and y_f.shape == y_b.shape and y_f.dtype == y_b.dtype
    B, T, C = y_f.shape
    out = torch.empty((B, T, 2 * C), device=y_f.device, dtype=y_f.dtype)
    prm = [t.detach().to(y_f.dtype).contiguous() for t in (ln_f.weight, ln_f.bias, ln_b.weight, ln_b.bias)]
    assert ln_f.eps == ln_b.eps
    with torch.cuda.device(y_f.device):
        _lib.check(_lib.lib().wkv6_b200_ln_pair_cat(_dtype_code(y_f), B * T, C, _p(y_f), _p(y_b), *[_p(t) for t in prm],
                                                    float(ln_f.eps), _p(out), _stream(y_f)))
    return out


def _tmix_fwd(x, x_edge, m, maa, reverse, nout):
    B, T, C = x.shape
    L, dt, st = _lib.lib(), _dtype_code(x), _stream(x)
    outs = [torch.empty_like(x) for _ in range(nout)]
    with torch.cuda.device(x.device):
        if m is None:
            _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, _p(x), _p(x_edge), None, _p(maa[0]), None, None, None, None, int(reverse),
                                            _p(outs[0]), None, None, None, None, st))
        else:
            _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, _p(x), _p(x_edge), _p(m), None, *[_p(t) for t in maa], int(reverse),
                                            None, *[_p(o) for o in outs], st))
    return outs


def _tmix_bwd(x, m, maa, grads, reverse):
    """gx, gm (or None), gmaa [NP, C] fp32 of one stage (include/wkv6_b200.h: wkv6_b200_tmix_mix_backward)."""
    B, T, C = x.shape
    L, dt, st = _lib.lib(), _dtype_code(x), _stream(x)
    npairs = 4 if m is not None else 1
    grads = [(torch.zeros_like(x) if g is None else g.contiguous()) for g in grads]
    gx = torch.empty_like(x)
    gm = torch.empty_like(m) if m is not None else None
    gmaa = torch.zeros((npairs, C), device=x.device, dtype=torch.float32)
    pad = [None] * (4 - npairs)
    with torch.cuda.device(x.device):
        _lib.check(L.wkv6_b200_tmix_mix_backward(dt, B, T, C, _p(x), _p(m), *[_p(t) for t in maa], *pad, *[_p(g) for g in grads], *pad,
                                                 int(reverse), _p(gx), _p(gm), _p(gmaa), st))
    return gx, gm, gmaa


def _flat(t, like):
    return t.detach().reshape(-1).to(like.dtype).contiguous()


class TmixStage1(torch.autograd.Function):
    """xxx = x + (shift(x) - x) * maa_x  (model.py:272-275) as one kernel, with its backward."""

    @staticmethod
    def forward(ctx, x, maa_x, reverse, x_edge):
        ctx.reverse, ctx.pshape, ctx.pdtype = bool(reverse), maa_x.shape, maa_x.dtype
        mx = _flat(maa_x, x)
        ctx.save_for_backward(x, mx)
        return _tmix_fwd(x, x_edge, None, [mx], reverse, 1)[0]

    @staticmethod
    def backward(ctx, g):
        x, mx = ctx.saved_tensors
        gx, _, gmaa = _tmix_bwd(x, None, [mx], [g], ctx.reverse)
        return gx, gmaa[0].to(ctx.pdtype).view(ctx.pshape), None, None


class TmixStage2(torch.autograd.Function):
    """x_c = x + (shift(x) - x) * (maa_c + m_c), c in r, k, v, w  (model.py:280-287) as one kernel, with its backward."""

    @staticmethod
    def forward(ctx, x, m, maa_r, maa_k, maa_v, maa_w, reverse, x_edge):
        ctx.reverse, ctx.pshape, ctx.pdtype = bool(reverse), maa_r.shape, maa_r.dtype
        maa = [_flat(t, x) for t in (maa_r, maa_k, maa_v, maa_w)]
        ctx.save_for_backward(x, m, *maa)
        return tuple(_tmix_fwd(x, x_edge, m, maa, reverse, 4))

    @staticmethod
    def backward(ctx, g_r, g_k, g_v, g_w):
        x, m, *maa = ctx.saved_tensors
        gx, gm, gmaa = _tmix_bwd(x, m, maa, [g_r, g_k, g_v, g_w], ctx.reverse)
        gp = [gmaa[i].to(ctx.pdtype).view(ctx.pshape) for i in range(4)]
        return gx, gm, gp[0], gp[1], gp[2], gp[3], None, None


def tmix_mix_autograd(x: torch.Tensor, maa_x, maa_rkvw, w1, w2, reverse: bool = False, x_edge: Optional[torch.Tensor] = None):
    """tmix_mix with gradients: the two elementwise stages are autograd Functions on the fused kernels (forward and
    backward), the LoRA GEMMs in between are torch ops.  Returns x_r, x_k, x_v, x_w."""
    assert x.is_cuda and x.dim() == 3
    x = x.contiguous()
    B, T, C = x.shape
    if x_edge is not None:
        x_edge = x_edge.detach().to(x.dtype).reshape(B, C).contiguous()
    xxx = TmixStage1.apply(x, maa_x, reverse, x_edge)
    m = torch.tanh(xxx.view(B * T, C) @ w1).view(B * T, 4, -1).transpose(0, 1)
    m = torch.bmm(m, w2).contiguous()                                    # [4, B*T, C]
    return TmixStage2.apply(x, m, *maa_rkvw, reverse, x_edge)


def tmix_mix(x: torch.Tensor, maa_x: torch.Tensor, maa_rkvw, w1: torch.Tensor, w2: torch.Tensor, reverse: bool = False,
             x_edge: Optional[torch.Tensor] = None):
    """Token shift + the five `maa` mixes of RWKV_Tmix_x060c.forward (model.py:270-289) with the two elementwise
    stages fused (one kernel each, bit-exact with the eager ops); the LoRA GEMMs in between stay library calls.
    x [B,T,C] contiguous; maa_* [1,1,C]; w1 [C,128], w2 [4,32,C]; x_edge [B,C] = the frame beyond the edge (streaming)
    or None (zeros).  Returns x_r, x_k, x_v, x_w.  Inference only (no autograd graph is recorded)."""
    assert x.is_cuda and x.is_contiguous() and x.dim() == 3
    B, T, C = x.shape
    L, dt, st = _lib.lib(), _dtype_code(x), _stream(x)
    flat = lambda t: t.detach().reshape(-1).contiguous()
    mx = flat(maa_x)
    mr, mk, mv, mw = (flat(t) for t in maa_rkvw)
    xxx = torch.empty_like(x)
    if x_edge is not None:
        x_edge = x_edge.detach().to(x.dtype).reshape(B, C).contiguous()
    with torch.cuda.device(x.device):
        _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, _p(x), _p(x_edge), None, _p(mx), None, None, None, None, int(bool(reverse)),
                                        _p(xxx), None, None, None, None, st))
        m = torch.tanh(xxx.view(B * T, C) @ w1.detach()).view(B * T, 4, -1).transpose(0, 1)
        m = torch.bmm(m, w2.detach()).contiguous()                      # [4, B*T, C]
        outs = [torch.empty_like(x) for _ in range(4)]
        _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, _p(x), _p(x_edge), _p(m), None, _p(mr), _p(mk), _p(mv), _p(mw), int(bool(reverse)),
                                        None, *[_p(o) for o in outs], st))
    return tuple(outs)
