"""Fused token-shift / maa-mix kernels of the time-mix prologue (csrc/wkv6_tmix.cu) against the eager torch
ops of the reference (restated in tests/_torch_ref.py from wenet/rwkv_v6/src/model.py:270-289), run on the CPU on the same
inputs.  The bar is BIT-EXACT: the kernels round wherever the eager ops materialise a tensor."""
import ctypes

import pytest
import torch

import _torch_ref as ref

pytestmark = pytest.mark.gpu

SHAPES = [(2, 250, 256), (1, 1, 64), (3, 2, 64), (2, 77, 512), (1, 1031, 1024)]


def _call(x, m, maa, reverse, x_edge=None):
    from paper_accurate_fast_cheap_b200 import _lib, ops
    B, T, C = x.shape
    L = _lib.lib()
    p = lambda t: None if t is None else ctypes.c_void_p(t.data_ptr())
    st = ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)
    dt = ops._dtype_code(x)
    n0 = _lib.launch_count()
    if m is None:
        out = torch.empty_like(x)
        _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, p(x), p(x_edge), None, p(maa[0]), None, None, None, None, int(reverse), p(out),
                                        None, None, None, None, st))
        res = (out,)
    else:
        outs = [torch.empty_like(x) for _ in range(4)]
        _lib.check(L.wkv6_b200_tmix_mix(dt, B, T, C, p(x), p(x_edge), p(m), None, p(maa[1]), p(maa[2]), p(maa[3]), p(maa[4]), int(reverse),
                                        None, *[p(o) for o in outs], st))
        res = tuple(outs)
    assert _lib.launch_count() - n0 == 1
    torch.cuda.synchronize()
    return res


@pytest.mark.parametrize("edge", [False, True], ids=["zero_edge", "edge_frame"])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("shape", SHAPES, ids=lambda s: "x".join(map(str, s)))
def test_tmix_mix_bit_exact(shape, reverse, dtype, edge):
    B, T, C = shape
    g = torch.Generator().manual_seed(B * 1000 + T + C)
    x = torch.randn(B, T, C, generator=g).to(dtype)
    xe = torch.randn(B, C, generator=g).to(dtype) if edge else None
    xed = xe.cuda() if edge else None
    m = (0.3 * torch.randn(4, B, T, C, generator=g)).to(dtype)
    maa = [torch.rand(1, 1, C, generator=g).to(dtype) for _ in range(5)]
    xd, md, maad = x.cuda(), m.cuda(), [t.cuda().reshape(-1).contiguous() for t in maa]
    (xxx,) = _call(xd, None, maad, reverse, xed)
    assert torch.equal(xxx.cpu(), ref.tmix_mix_stage1(x, maa[0], reverse, xe))
    outs = _call(xd, md, maad, reverse, xed)
    want = ref.tmix_mix_stage2(x, m, maa[1], maa[2], maa[3], maa[4], reverse, xe)
    for name, a, b in zip("rkvw", outs, want):
        assert torch.equal(a.cpu(), b), name


def test_tmix_mix_rejects_bad_arguments():
    from paper_accurate_fast_cheap_b200 import _lib
    x = torch.zeros(1, 4, 60, device="cuda", dtype=torch.bfloat16)          # C % 8 != 0
    with pytest.raises(_lib.Wkv6Error):
        _call(x, None, [torch.zeros(60, device="cuda", dtype=torch.bfloat16)] * 5, False)


@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_prologue_fused_equals_eager_on_gpu(dtype, reverse):
    """RWKV_Tmix_x060c.prologue under no_grad (fused stages) returns the very bits of the eager path (same GEMMs)."""
    from paper_accurate_fast_cheap_b200 import wrappers
    torch.manual_seed(5)
    blk = wrappers.RWKV_Tmix_x060c(64, 12, 256, 256, 3).to(device="cuda", dtype=dtype)
    with torch.no_grad():
        blk.time_maa_rkvw_w1.normal_(0, 0.05)
        blk.time_decay_w1.normal_(0, 0.05)
    x = torch.randn(2, 93, 256, device="cuda").to(dtype)
    with torch.no_grad():
        fused = blk.prologue(x, reverse)
    eager = blk.prologue(x, reverse)                                        # grad mode: eager torch ops
    for name, a, b in zip("rkvw", fused, eager):
        assert torch.equal(a, b.detach()), name
