"""Pins the C oracle (oracle/) against torch.autograd of a plain fp64 recurrence and checks the
literal restatements of the reference backward kernels against the derivation-based one."""
import pytest
import torch

import oracle
from _torch_ref import make_inputs, wkv6_recurrence


def _autograd(inp, rev=False, with_state=False):
    leaves = {n: inp[n].clone().requires_grad_(True) for n in ("r", "k", "v", "w", "u")}
    s0 = inp["s0"].clone().requires_grad_(True) if with_state else None
    y, sT = wkv6_recurrence(leaves["r"], leaves["k"], leaves["v"], leaves["w"], leaves["u"], s0, rev)
    loss = (y * inp["gy"]).sum()
    if with_state:
        loss = loss + (sT * inp["gsT"]).sum()
    loss.backward()
    g = {n: leaves[n].grad for n in leaves}
    if with_state:
        g["s0"] = s0.grad
    return y.detach(), sT.detach(), g


@pytest.mark.parametrize("rev", [False, True])
@pytest.mark.parametrize("with_state", [False, True])
def test_fwd_bwd_f64_matches_autograd(rev, with_state):
    B, T, H, N = 2, 37, 3, 8
    inp = make_inputs(B, T, H, N, with_state=with_state)
    y_ref, sT_ref, g = _autograd(inp, rev, with_state)
    s0 = inp.get("s0")
    y, sT = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], s0, rev=rev,
                       want_state=True, prec="f64")
    assert oracle.rel_err(y, y_ref) < 1e-13
    assert oracle.rel_err(sT, sT_ref) < 1e-13
    gr, gk, gv, gw, gu, gs0 = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"],
                                         s0=s0, gsT=inp.get("gsT"), rev=rev, prec="f64")
    for name, got, ref in (("gr", gr, g["r"]), ("gk", gk, g["k"]), ("gv", gv, g["v"]), ("gw", gw, g["w"]),
                           ("gu", gu.sum(0).view(H, N), g["u"])):
        assert oracle.rel_err(got, ref) < 1e-12, name
    if with_state:
        assert oracle.rel_err(gs0, g["s0"]) < 1e-12


@pytest.mark.parametrize("prec,tol", [("f64", 1e-12), ("f32", 2e-5)])
def test_literal_bwd_matches_autograd(prec, tol):
    """kernel_backward_101/102/103/201 restated literally == autograd of the recurrence (S0 = 0)."""
    B, T, H, N = 2, 50, 2, 16
    inp = make_inputs(B, T, H, N)
    _, _, g = _autograd(inp)
    gr, gk, gv, gw, gu = oracle.bwd_ref(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec=prec)
    for name, got, ref in (("gr", gr, g["r"]), ("gk", gk, g["k"]), ("gv", gv, g["v"]), ("gw", gw, g["w"]),
                           ("gu", gu.sum(0).view(H, N), g["u"])):
        assert oracle.rel_err(got, ref) < tol, name
    # the reference hard-codes these (wkv6_cuda.cu:237,262); exact for S0 = 0
    assert float(gw[:, 0].abs().max()) == 0.0 and float(gw[:, -1].abs().max()) == 0.0


@pytest.mark.parametrize("prec,tol", [("f64", 1e-12), ("f32", 2e-5)])
def test_literal_state_bwd_matches_autograd(prec, tol):
    """kernel_backward_111/222 (wkv6state_cuda.cu) restated literally == autograd with S0 != 0."""
    B, T, H, N = 2, 41, 2, 16
    inp = make_inputs(B, T, H, N, with_state=True)
    inp["gsT"] = torch.zeros_like(inp["gsT"])          # the reference has no final-state output
    _, _, g = _autograd(inp, with_state=True)
    gr, gk, gv, gw, gu, gs = oracle.bwd_state_ref(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["s0"],
                                                  inp["gy"], prec=prec)
    for name, got, ref in (("gr", gr, g["r"]), ("gk", gk, g["k"]), ("gv", gv, g["v"]), ("gw", gw, g["w"]),
                           ("gu", gu.sum(0).view(H, N), g["u"]), ("gs", gs, g["s0"])):
        assert oracle.rel_err(got, ref) < tol, name


def test_f32_forward_close_to_f64_and_stress():
    for mode in ("init", "stress"):
        inp = make_inputs(2, 300, 4, 64, w_mode=mode, dtype=torch.float32)
        y64, s64 = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], want_state=True, prec="f64")
        y32, s32 = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], want_state=True, prec="f32")
        assert oracle.rel_err(y32, y64) < 1e-5
        assert oracle.rel_err(s32, s64) < 1e-5


def test_flip_free_reverse_equals_flip():
    """running t = T-1..0 in place == flip -> scan -> flip (bit exact)."""
    inp = make_inputs(2, 33, 2, 16, dtype=torch.float32)
    f = lambda x: torch.flip(x, [1]).contiguous()
    y_rev, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], rev=True)
    y_flip, _ = oracle.fwd(f(inp["r"]), f(inp["k"]), f(inp["v"]), f(inp["w"]), inp["u"])
    assert torch.equal(y_rev, f(y_flip))


def test_split_run_state_carry():
    """Two calls chained through S_T == one call over the concatenation (bit exact in f32)."""
    inp = make_inputs(2, 64, 2, 16, dtype=torch.float32)
    a = {n: inp[n][:, :40] for n in "rkvw"}
    b = {n: inp[n][:, 40:] for n in "rkvw"}
    y, sT = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], want_state=True)
    y1, s1 = oracle.fwd(a["r"], a["k"], a["v"], a["w"], inp["u"], want_state=True)
    y2, s2 = oracle.fwd(b["r"], b["k"], b["v"], b["w"], inp["u"], s0=s1, want_state=True)
    assert torch.equal(torch.cat([y1, y2], 1), y) and torch.equal(s2, sT)
