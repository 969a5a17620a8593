"""GPU parity of the tcgen05 chunk kernel (bf16 forward): same tolerance as the exact scan (2e-2 max-relative vs the
fp64 oracle on the same rounded inputs) for y and the final state, plus agreement with the exact scan kernel."""
import pytest
import torch

import oracle
from test_gpu_parity import DEV, _cuda, _inputs

pytestmark = pytest.mark.gpu


@pytest.fixture()
def chunk_algo():
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import _lib
    pk.set_algo(_lib.ALGO_CHUNK)
    yield pk
    pk.set_algo(_lib.ALGO_AUTO)


@pytest.mark.parametrize("rev", [False, True])
@pytest.mark.parametrize("B,T,H,with_state,w_mode", [
    (1, 64, 1, False, "init"),       # exactly one chunk
    (2, 250, 4, False, "init"),      # BASELINE config 1 shape (bf16)
    (1, 1, 1, False, "init"),
    (2, 130, 2, True, "init"),       # ragged tail + initial state
    (3, 53, 2, True, "stress"),      # strong decays: exercises the sub-block factorisation and the clamp
    (2, 750, 8, False, "init"),      # BASELINE config 2 sequence length
])
def test_chunk_forward_matches_oracle(chunk_algo, rev, B, T, H, with_state, w_mode):
    pk = chunk_algo
    inp = _inputs(B, T, H, torch.bfloat16, w_mode, with_state)
    g = _cuda(inp)
    n0 = pk.launch_count()
    y, sT = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], s0=g.get("s0"), reverse=rev, want_state=True)
    assert pk.launch_count() - n0 == 1
    y_ref, sT_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp.get("s0"), rev=rev,
                               want_state=True, prec="f64")
    assert oracle.rel_err(y, y_ref) <= 2e-2
    assert oracle.rel_err(sT, sT_ref) <= 2e-2


def test_chunk_bidirectional_one_launch(chunk_algo):
    pk = chunk_algo
    a = _inputs(2, 200, 4, torch.bfloat16, seed=1)
    b = _inputs(2, 200, 4, torch.bfloat16, seed=2, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    n0 = pk.launch_count()
    y_f, y_b, _, _ = pk.bi_forward_raw(tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu"))
    assert pk.launch_count() - n0 == 1
    ya, _ = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], prec="f64")
    yb, _ = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], rev=True, prec="f64")
    assert oracle.rel_err(y_f, ya) <= 2e-2 and oracle.rel_err(y_b, yb) <= 2e-2


@pytest.mark.parametrize("algo", ["scan", "chunk"])
@pytest.mark.parametrize("rev", [False, True])
@pytest.mark.parametrize("B,T,H,with_state,w_mode", [
    (1, 64, 1, False, "init"),
    (2, 130, 2, True, "init"),       # ragged tail, initial state, gradient into the final state
    (3, 53, 2, True, "stress"),
    (2, 750, 8, False, "init"),      # BASELINE config 2 sequence length
])
def test_bf16_backward_both_kernels(algo, rev, B, T, H, with_state, w_mode):
    """bf16 backward on the exact scan kernels and on the tcgen05 chunk kernel (checkpoints recomputed internally)."""
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import _lib
    pk.set_algo(_lib.ALGO_SCAN if algo == "scan" else _lib.ALGO_CHUNK)
    try:
        inp = _inputs(B, T, H, torch.bfloat16, w_mode, with_state)
        g = _cuda(inp)
        n0 = pk.launch_count()
        out = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"], s0=g.get("s0"), gsT=g.get("gsT"),
                                   reverse=rev, want_gs0=with_state)
        assert pk.launch_count() - n0 == 2      # scan: sweeps A + B; chunk: checkpoint recompute + one sweep
    finally:
        pk.set_algo(_lib.ALGO_AUTO)
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], s0=inp.get("s0"), gsT=inp.get("gsT"),
                     rev=rev, prec="f64")
    for name, a, b in zip(("gr", "gk", "gv", "gw", "gu"), out[:5], ref[:5]):
        assert oracle.rel_err(a, b) <= 2e-2, name
    if with_state:
        assert oracle.rel_err(out[5], ref[5]) <= 2e-2


def test_chunk_backward_uses_forward_checkpoints():
    """Training path: the forward saves the chunk-start states, the backward is ONE launch for both directions."""
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import ops
    a = _inputs(2, 300, 4, torch.bfloat16, seed=1)
    b = _inputs(2, 300, 4, torch.bfloat16, seed=2, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    ck = ops.checkpoint_buffer(ga["r"], 2)
    assert ck is not None and ck.shape[1] == 2 * 4 * 5 * 64 * 64 * 2
    n0 = pk.launch_count()
    y_f, y_b, _, _ = pk.bi_forward_raw(f, r_, ckpt=ck)
    of, ob = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
    assert pk.launch_count() - n0 == 2
    of2, ob2 = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"])          # recompute path: identical results
    for x, y in zip(of[:5] + ob[:5], of2[:5] + ob2[:5]):
        assert torch.equal(x, y)
    ra = oracle.bwd(a["r"], a["k"], a["v"], a["w"], a["u"], a["gy"], prec="f64")
    rb = oracle.bwd(b["r"], b["k"], b["v"], b["w"], b["u"], b["gy"], rev=True, prec="f64")
    for got, want in ((of, ra), (ob, rb)):
        for name, x, y in zip(("gr", "gk", "gv", "gw", "gu"), got[:5], want[:5]):
            assert oracle.rel_err(x, y) <= 2e-2, name
    # autograd: BiWKV6 keeps the checkpoints between forward and backward
    lf = [t.clone().requires_grad_(True) for t in f]
    lb = [t.clone().requires_grad_(True) for t in r_]
    n0 = pk.launch_count()
    yf, yb = pk.bi_wkv6(lf, lb)
    ((yf.float() * ga["gy"].float()).sum() + (yb.float() * gb["gy"].float()).sum()).backward()
    assert pk.launch_count() - n0 == 2
    assert oracle.rel_err(lf[3].grad, ra[3]) <= 2e-2 and oracle.rel_err(lb[3].grad, rb[3]) <= 2e-2
