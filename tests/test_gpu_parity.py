"""GPU parity: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
Tolerance (SURVEY 8c / BASELINE north_star): max|a-b|/max|b| <= 1e-5 for fp32 I/O, <= 2e-2 for bf16 I/O,
for y, the final state and every gradient, measured against the fp64 oracle on the same rounded inputs."""
import pytest
import torch

import oracle
from _torch_ref import make_inputs

pytestmark = pytest.mark.gpu

TOL = {torch.float32: 1e-5, torch.bfloat16: 2e-2}
DEV = "cuda:0"


def _pkg():
    import paper_accurate_fast_cheap_b200 as pk
    return pk


def _inputs(B, T, H, dtype, w_mode="init", with_state=False, seed=777, layer=3):
    inp = make_inputs(B, T, H, 64, seed=seed, dtype=torch.float64, w_mode=w_mode, with_state=with_state, layer=layer)
    for n in ("r", "k", "v", "w", "u", "gy"):
        inp[n] = inp[n].to(dtype)            # round ONCE; oracle and kernel see identical values
    if with_state:
        inp["s0"] = inp["s0"].float()
        inp["gsT"] = inp["gsT"].float()
    return inp


def _cuda(inp):
    return {n: t.to(DEV).contiguous() for n, t in inp.items()}


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rev", [False, True])
@pytest.mark.parametrize("B,T,H,with_state,w_mode", [
    (2, 250, 4, False, "init"),      # BASELINE config 1 shape
    (1, 1, 1, False, "init"),        # single step
    (2, 16, 2, True, "init"),        # exactly one tile
    (3, 53, 2, True, "stress"),      # ragged tail, strong decays
    (1, 300, 8, True, "init"),
])
def test_forward_matches_oracle(dtype, rev, B, T, H, with_state, w_mode):
    pk = _pkg()
    inp = _inputs(B, T, H, dtype, w_mode, with_state)
    g = _cuda(inp)
    y, sT = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], s0=g.get("s0"), reverse=rev, want_state=True)
    y_ref, sT_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp.get("s0"), rev=rev,
                               want_state=True, prec="f64")
    tol = TOL[dtype]
    assert y.dtype == dtype and sT.dtype == torch.float32
    assert oracle.rel_err(y, y_ref) <= tol
    assert oracle.rel_err(sT, sT_ref) <= tol   # bf16: the chunk kernel feeds bf16 operands to the tensor cores


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("rev", [False, True])
@pytest.mark.parametrize("B,T,H,with_state,w_mode", [
    (2, 250, 4, False, "init"),
    (1, 2, 1, False, "init"),
    (2, 33, 2, True, "init"),
    (2, 70, 2, True, "stress"),
])
def test_backward_matches_oracle(dtype, rev, B, T, H, with_state, w_mode):
    pk = _pkg()
    inp = _inputs(B, T, H, dtype, w_mode, with_state)
    g = _cuda(inp)
    gr, gk, gv, gw, gu, gs0 = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"], s0=g.get("s0"),
                                                   gsT=g.get("gsT"), reverse=rev, want_gs0=with_state)
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], s0=inp.get("s0"),
                     gsT=inp.get("gsT"), rev=rev, prec="f64")
    tol = TOL[dtype]
    for name, got, want in zip(("gr", "gk", "gv", "gw", "gu"), (gr, gk, gv, gw, gu), ref[:5]):
        assert got.dtype == dtype
        assert oracle.rel_err(got, want) <= tol, name
    if with_state:
        assert oracle.rel_err(gs0, ref[5]) <= 1e-5


def test_backward_matches_literal_reference_restatement():
    """S0 = 0: against the literal restatement of kernel_backward_101/102/103/201, incl. gw[0] = gw[T-1] = 0."""
    pk = _pkg()
    inp = _inputs(2, 120, 2, torch.float32)
    g = _cuda(inp)
    got = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"])
    ref = oracle.bwd_ref(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    for name, a, b in zip(("gr", "gk", "gv", "gw", "gu"), got[:5], ref):
        assert oracle.rel_err(a, b) <= 1e-5, name
    assert float(got[3][:, 0].abs().max()) == 0.0 and float(got[3][:, -1].abs().max()) == 0.0


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_bidirectional_one_launch_and_direction_flags(dtype):
    pk = _pkg()
    from paper_accurate_fast_cheap_b200 import _lib
    a = _inputs(2, 77, 4, dtype, seed=1)
    b = _inputs(2, 77, 4, dtype, seed=2, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f = tuple(ga[n] for n in "rkvwu")
    r_ = tuple(gb[n] for n in "rkvwu")
    tol = TOL[dtype]
    ya_ref, _ = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], prec="f64")
    yb_ref, _ = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], rev=True, prec="f64")
    n0 = pk.launch_count()
    y_f, y_b, _, _ = pk.bi_forward_raw(f, r_)
    assert pk.launch_count() - n0 == 1
    assert oracle.rel_err(y_f, ya_ref) <= tol and oracle.rel_err(y_b, yb_ref) <= tol
    # flip-free reversed direction == flip -> scan -> flip
    fl = lambda t: torch.flip(t, [1]).contiguous()
    y_flip, _ = pk.wkv6_forward_raw(fl(gb["r"]), fl(gb["k"]), fl(gb["v"]), fl(gb["w"]), gb["u"])
    if dtype == torch.float32:
        assert torch.equal(fl(y_flip), y_b)          # exact scan: bit-identical
    else:
        assert oracle.rel_err(fl(y_flip), y_b) <= 2e-2  # chunk kernel: chunk boundaries differ after a flip
    # direction dropout flags
    y_f2, y_b2, _, _ = pk.bi_forward_raw(f, r_, flags=_lib.BI_L2R)
    assert y_b2 is None and torch.equal(y_f2, y_f)
    y_f3, y_b3, _, _ = pk.bi_forward_raw(f, r_, flags=_lib.BI_R2L)
    assert y_f3 is None and torch.equal(y_b3, y_b)
    # backward of both directions
    of, ob = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"])
    ra = oracle.bwd(a["r"], a["k"], a["v"], a["w"], a["u"], a["gy"], prec="f64")
    rb = oracle.bwd(b["r"], b["k"], b["v"], b["w"], b["u"], b["gy"], rev=True, prec="f64")
    for got, want in ((of, ra), (ob, rb)):
        for name, x, y in zip(("gr", "gk", "gv", "gw", "gu"), got[:5], want[:5]):
            assert oracle.rel_err(x, y) <= tol, name


def test_split_run_state_carry_and_long_sequence():
    """Size-independent property at a long T: two chained calls == one call (chunked state carry)."""
    pk = _pkg()
    inp = _inputs(1, 5000, 2, torch.float32)
    g = _cuda(inp)
    y, sT = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], want_state=True)
    cut = 2077
    a = {n: g[n][:, :cut].contiguous() for n in "rkvw"}
    b = {n: g[n][:, cut:].contiguous() for n in "rkvw"}
    y1, s1 = pk.wkv6_forward_raw(a["r"], a["k"], a["v"], a["w"], g["u"], want_state=True)
    y2, s2 = pk.wkv6_forward_raw(b["r"], b["k"], b["v"], b["w"], g["u"], s0=s1, want_state=True)
    assert oracle.rel_err(torch.cat([y1, y2], 1), y) <= 1e-6
    assert oracle.rel_err(s2, sT) <= 1e-6
    y_ref, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], prec="f64")
    assert oracle.rel_err(y, y_ref) <= 1e-5


def test_autograd_functions_and_legacy_ops():
    pk = _pkg()
    pk.load_torch_ops()
    inp = _inputs(2, 40, 2, torch.bfloat16)
    g = _cuda(inp)
    leaves = {n: g[n].clone().requires_grad_(True) for n in "rkvwu"}
    y = pk.RUN_CUDA_RWKV6(*(leaves[n] for n in "rkvwu"))
    y.backward(g["gy"])
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    assert oracle.rel_err(leaves["u"].grad, ref[4].sum(0).view(2, 64)) <= 2e-2
    assert oracle.rel_err(leaves["w"].grad, ref[3]) <= 2e-2
    # the reference's own dispatcher entry points (wkv6_op.cpp:34-41)
    y2 = torch.empty_like(g["r"])
    torch.ops.wkv6.forward(2, 40, 128, 2, g["r"], g["k"], g["v"], g["w"], g["u"], y2)
    assert torch.equal(y2, y.detach())
    f32 = {n: g[n].float() for n in ("r", "k", "v", "w", "u", "gy")}
    y3 = torch.empty_like(f32["r"])
    torch.ops.wkv6.forward_fp32(2, 40, 128, 2, f32["r"], f32["k"], f32["v"], f32["w"], f32["u"], y3)
    y_ref, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], prec="f64")
    assert oracle.rel_err(y3, y_ref) <= 1e-5
    outs = [torch.empty_like(f32["r"]) for _ in range(4)] + [torch.empty(2, 128, device=DEV)]
    torch.ops.wkv6.backward_fp32(2, 40, 128, 2, f32["r"], f32["k"], f32["v"], f32["w"], f32["u"], f32["gy"], *outs)
    for name, a, b in zip(("gr", "gk", "gv", "gw", "gu"), outs, ref[:5]):
        assert oracle.rel_err(a, b) <= 1e-5, name
    # state-carrying autograd op: gradient w.r.t. S_0 and through S_T
    s = _inputs(1, 30, 2, torch.float32, with_state=True)
    gs = _cuda(s)
    lv = {n: gs[n].clone().requires_grad_(True) for n in ("r", "k", "v", "w", "u", "s0")}
    yy, sT = pk.wkv6(lv["r"], lv["k"], lv["v"], lv["w"], lv["u"], lv["s0"])
    ((yy * gs["gy"]).sum() + (sT * gs["gsT"]).sum()).backward()
    rs = oracle.bwd(s["r"], s["k"], s["v"], s["w"], s["u"], s["gy"], s0=s["s0"], gsT=s["gsT"], prec="f64")
    assert oracle.rel_err(lv["s0"].grad, rs[5]) <= 1e-5
    assert oracle.rel_err(lv["w"].grad, rs[3]) <= 1e-5


def test_rejects_bad_arguments():
    pk = _pkg()
    x = torch.zeros(1, 4, 96, device=DEV)
    u = torch.zeros(1, 96, device=DEV)
    with pytest.raises(Exception):
        pk.wkv6_forward_raw(x, x, x, x, u)            # head size != 64
    c = torch.zeros(1, 4, 64)
    with pytest.raises(AssertionError):
        pk.wkv6_forward_raw(c, c, c, c, torch.zeros(1, 64))   # CPU tensors: no fallback


def _golden_cases():
    import glob
    import os
    return sorted(glob.glob(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "refcuda_*.npz")))


@pytest.mark.parametrize("path", _golden_cases(), ids=[p.split("refcuda_")[-1][:-4] for p in _golden_cases()])
def test_matches_reference_cuda_golden_vectors(path):
    """CUDA path vs the committed outputs of the reference's own CUDA kernels (tests/golden/refcuda_*.npz).
    Both sides carry their own rounding: the bound is the sum of the two errors vs fp64 truth."""
    from test_golden import load_case
    pk = _pkg()
    c = load_case(path)
    g = _cuda(c["inp"])
    bf = c["dtype"] == "torch.bfloat16"
    tol = 2e-2 if bf else 1e-5
    y, _ = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"])
    assert oracle.rel_err(y[:, ::c["stride"]], c["outs"]["y"]) <= tol
    if c["bwd"]:
        got = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"])
        for name, a in zip(("gr", "gk", "gv", "gw", "gu"), got[:5]):
            assert oracle.rel_err(a, c["outs"][name]) <= tol, name


def test_scriptmodule_in_the_reference_style_runs_on_these_kernels():
    """INTEGRATION level 1, run rather than asserted: a TorchScript ScriptModule written the way the reference's model.py is
    (MyModule = torch.jit.ScriptModule, model.py:33-35; the WKV_6 autograd.Function calling torch.ops.wkv6.forward /
    backward positionally on caller-allocated outputs, model.py:108-159) dispatches into libwkv6_b200.so after
    install_ops(), forward and backward, bf16, on the GPU."""
    pk = _pkg()
    from paper_accurate_fast_cheap_b200 import wrappers
    wrappers.install_ops()

    class WKV_6_ref(torch.autograd.Function):            # the reference's function, restated with its call sites
        @staticmethod
        def forward(ctx, B: int, T: int, C: int, H: int, r, k, v, w, u):
            ctx.B, ctx.T, ctx.C, ctx.H = B, T, C, H
            ctx.save_for_backward(r, k, v, w, u)
            y = torch.empty((B, T, C), device=r.device, dtype=torch.bfloat16, memory_format=torch.contiguous_format)
            torch.ops.wkv6.forward(B, T, C, H, r, k, v, w, u, y)
            return y

        @staticmethod
        def backward(ctx, gy):
            B, T, C, H = ctx.B, ctx.T, ctx.C, ctx.H
            r, k, v, w, u = ctx.saved_tensors
            mk = lambda *s: torch.empty(s, device=gy.device, requires_grad=False, dtype=torch.bfloat16)
            gr, gk, gv, gw, gu = mk(B, T, C), mk(B, T, C), mk(B, T, C), mk(B, T, C), mk(B, C)
            torch.ops.wkv6.backward(B, T, C, H, r, k, v, w, u, gy.contiguous(), gr, gk, gv, gw, gu)
            return (None, None, None, None, gr, gk, gv, gw, torch.sum(gu, 0).view(H, C // H))

    class Block(torch.jit.ScriptModule):
        def __init__(self, C: int, H: int):
            super().__init__()
            self.C, self.H = C, H
            self.receptance = torch.nn.Linear(C, C, bias=False)

        @torch.jit.script_method
        def mix(self, x):
            return self.receptance(x)

        def forward(self, x, k, v, w, u):
            r = self.mix(x)
            return WKV_6_ref.apply(x.size(0), x.size(1), self.C, self.H, r.contiguous(), k, v, w, u)

    inp = _inputs(2, 130, 2, torch.bfloat16)
    g = _cuda(inp)
    blk = Block(128, 2).to(DEV).to(torch.bfloat16)
    with torch.no_grad():
        blk.receptance.weight.copy_(torch.eye(128))       # r = x exactly: the oracle can check the result
    leaves = {n: g[n].clone().requires_grad_(True) for n in "rkvwu"}
    n0 = pk.launch_count()
    y = blk(leaves["r"], leaves["k"], leaves["v"], leaves["w"], leaves["u"])
    y.backward(g["gy"])
    assert pk.launch_count() - n0 == 3                    # forward + (checkpoint recompute, sweep): the legacy ops keep no state
    y_ref, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], prec="f64")
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    assert oracle.rel_err(y, y_ref) <= 2e-2
    for n, want in zip("kvw", ref[1:4]):
        assert oracle.rel_err(leaves[n].grad, want) <= 2e-2, n
    assert oracle.rel_err(leaves["u"].grad, ref[4].sum(0).view(2, 64)) <= 2e-2
