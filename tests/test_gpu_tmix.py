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


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("reverse", [False, True])
@pytest.mark.parametrize("shape", [(2, 250, 256), (1, 1, 64), (3, 2, 64), (2, 77, 512), (1, 1031, 1024), (16, 750, 512)],
                         ids=lambda s: "x".join(map(str, s)))
def test_tmix_mix_backward_matches_autograd(shape, reverse, dtype):
    """Backward of the two fused stages (ops.TmixStage1 / TmixStage2 -> wkv6_b200_tmix_mix_backward) against torch.autograd
    of the eager ops in float64 on the same (rounded) inputs: gradients w.r.t. x, m and the five maa vectors."""
    import oracle
    from paper_accurate_fast_cheap_b200 import ops
    B, T, C = shape
    g = torch.Generator().manual_seed(B * 1000 + T + C)
    x = torch.randn(B, T, C, generator=g).to(dtype)
    m = (0.3 * torch.randn(4, B * T, C, generator=g)).to(dtype)
    maa = [torch.rand(1, 1, C, generator=g).to(dtype) for _ in range(5)]
    gout = [torch.randn(B, T, C, generator=g).to(dtype) for _ in range(5)]
    # reference: float64 autograd of the eager formulas
    xr = x.double().requires_grad_(True)
    mr = m.double().view(4, B, T, C).requires_grad_(True)
    maar = [t.double().requires_grad_(True) for t in maa]
    xxx = ref.tmix_mix_stage1(xr, maar[0], reverse)
    outs = ref.tmix_mix_stage2(xr, mr, maar[1], maar[2], maar[3], maar[4], reverse)
    loss = (xxx * gout[0].double()).sum() + sum((o * go.double()).sum() for o, go in zip(outs, gout[1:]))
    loss.backward()
    # kernels
    xd = x.cuda().requires_grad_(True)
    md = m.cuda().requires_grad_(True)
    maad = [t.cuda().requires_grad_(True) for t in maa]
    n0 = ops._lib.launch_count()
    xxx_d = ops.TmixStage1.apply(xd, maad[0], reverse, None)
    outs_d = ops.TmixStage2.apply(xd, md, maad[1], maad[2], maad[3], maad[4], reverse, None)
    loss_d = (xxx_d.float() * gout[0].cuda().float()).sum() + sum((o.float() * go.cuda().float()).sum() for o, go in zip(outs_d, gout[1:]))
    loss_d.backward()
    assert ops._lib.launch_count() - n0 == 4            # two forward and two backward launches
    tol = 2e-6 if dtype == torch.float32 else 1e-2
    assert oracle.rel_err(xd.grad, xr.grad) <= tol
    assert oracle.rel_err(md.grad.view(4, B, T, C), mr.grad) <= tol
    for i in range(5):
        # column sums over B*T rows: fp32 accumulation (atomics), then one rounding to the parameter dtype
        assert oracle.rel_err(maad[i].grad, maar[i].grad) <= (2e-5 if dtype == torch.float32 else 1e-2), i


def test_prologue_trains_through_the_fused_stages():
    """RWKV_Tmix_x060c.prologue with autograd on uses the fused stages (it fell back to ~14 eager kernels per
    direction) and its parameter / input gradients agree with the eager formulation."""
    import oracle
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import ops
    torch.manual_seed(3)
    blk = pk.RWKV_Tmix_x060c(64, 12, 256, 256, 4).cuda()
    with torch.no_grad():
        blk.time_maa_rkvw_w1.normal_(0, 0.05)
        blk.time_decay_w1.normal_(0, 0.05)
    x = torch.randn(2, 90, 256, device="cuda")
    gy = [torch.randn(2, 90, 256, device="cuda") for _ in range(4)]
    import copy
    blk64 = copy.deepcopy(blk).double()                  # float64 has no fused kernels: prologue() takes its eager branch
    for reverse in (False, True):
        res = []
        for fused in (True, False):
            mod = blk if fused else blk64
            mod.zero_grad(set_to_none=True)
            xi = (x if fused else x.double()).clone().requires_grad_(True)
            n0 = pk.launch_count()
            outs = mod.prologue(xi, reverse)
            assert (pk.launch_count() - n0 == 2) == fused
            sum((o.double() * g_.double()).sum() for o, g_ in zip(outs, gy)).backward()
            res.append([xi.grad.double()] + [p.grad.double() for p in mod.parameters() if p.grad is not None])
        assert len(res[0]) == len(res[1]) and len(res[0]) >= 12
        for a, b in zip(*res):
            assert oracle.rel_err(a, b) <= 5e-5


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rows,C", [(1, 64), (77, 256), (12000, 512), (5, 1024), (3, 2048)])
def test_ln_pair_cat_matches_layer_norm(rows, C, dtype):
    """The pair of ln_x of a bidirectional block's epilogue written side by side (ops.ln_pair_cat) against F.layer_norm."""
    import oracle
    from paper_accurate_fast_cheap_b200 import ops
    g = torch.Generator().manual_seed(rows + C)
    y = [(2.0 * torch.randn(1, rows, C, generator=g) + 0.5).to(dtype).cuda() for _ in range(2)]
    lns = []
    for _ in range(2):
        ln = torch.nn.LayerNorm(C).to(dtype).cuda()
        with torch.no_grad():
            ln.weight.copy_((1 + 0.2 * torch.randn(C, generator=g)).to(dtype))
            ln.bias.copy_((0.2 * torch.randn(C, generator=g)).to(dtype))
        lns.append(ln)
    n0 = ops._lib.launch_count()
    z = ops.ln_pair_cat(y[0], y[1], lns[0], lns[1])
    assert ops._lib.launch_count() - n0 == 1 and tuple(z.shape) == (1, rows, 2 * C) and z.dtype == dtype
    for d in range(2):
        want = torch.nn.functional.layer_norm(y[d].double(), (C,), lns[d].weight.double(), lns[d].bias.double(), lns[d].eps)
        assert oracle.rel_err(z[..., d * C:(d + 1) * C], want) <= (2e-6 if dtype == torch.float32 else 8e-3)
