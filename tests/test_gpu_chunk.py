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
