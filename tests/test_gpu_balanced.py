"""GPU parity of the balanced persistent grid of the bf16 chunk backward (more sequences than SMs: DESIGN.md 4.2).
With the scratch from wkv6_b200_backward_scratch_bytes every direction's (sequence, chunk) units are cut into equal
contiguous ranges, one CTA per range; a sequence that is cut between two CTAs hands its reverse state and partial gu over
through a record in the scratch.  Checked against the one-CTA-per-sequence launch of the same kernel (bit-equal except
gu, whose partial sums are regrouped) and against the fp64 oracle."""
import pytest
import torch

import oracle
from test_gpu_parity import TOL, _cuda, _inputs

pytestmark = pytest.mark.gpu


def _pk():
    import paper_accurate_fast_cheap_b200 as pk
    return pk


def _run(B, T, H, flags, with_state, seed=61):
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import _lib, ops
    a = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=seed)
    b = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=seed + 1, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    run_f, run_b = bool(flags & 1), bool(flags & 2)
    ndir = int(run_f) + int(run_b)
    nb = _lib.lib().wkv6_b200_backward_scratch_bytes(0, B, T, H * 64, H, ndir)
    assert nb > 0, "this shape should take the balanced grid"
    ck = ops.checkpoint_buffer(ga["r"], 2)
    pk.bi_forward_raw(f, r_, flags=flags, s0_f=ga.get("s0"), s0_b=gb.get("s0"), ckpt=ck)
    kw = dict(flags=flags, s0_f=ga.get("s0"), s0_b=gb.get("s0"), gsT_f=ga.get("gsT"), gsT_b=gb.get("gsT"),
              want_gs0=with_state, ckpt=ck)
    gyf, gyb = (ga["gy"] if run_f else None), (gb["gy"] if run_b else None)
    ops.set_balanced(False)
    try:
        of0, ob0 = pk.bi_backward_raw(f, r_, gyf, gyb, **kw)
    finally:
        ops.set_balanced(True)
    n0 = pk.launch_count()
    of, ob = pk.bi_backward_raw(f, r_, gyf, gyb, **kw)
    of2, ob2 = pk.bi_backward_raw(f, r_, gyf, gyb, **kw)       # again: the consumed flags were reset
    torch.cuda.synchronize()
    assert pk.launch_count() - n0 == 2, "one launch per call"
    return (a, b), (of0, ob0), (of, ob), (of2, ob2), (run_f, run_b)


@pytest.mark.parametrize("B,T,H,flags,with_state", [
    (16, 750, 8, 3, False),      # the bench shape (configs[1]): 256 sequences x 12 chunks on 2 x 74 CTAs
    (20, 333, 8, 3, True),       # ragged T, initial state, gradient into the final state, gs0
    (19, 200, 8, 1, False),      # one direction dropped: 152 sequences on 148 CTAs (4.1 chunks per CTA, every sequence cut)
    (10, 64, 16, 2, True),       # one chunk per sequence: 160 units, nothing to cut but the persistent walk
    (40, 130, 4, 3, False),      # 3 chunks per sequence, 160 sequences per direction
])
def test_balanced_grid_matches_one_cta_per_sequence(B, T, H, flags, with_state):
    (a, b), ref, got, got2, runs = _run(B, T, H, flags, with_state)
    nout = 6 if with_state else 5
    names = ("gr", "gk", "gv", "gw", "gu", "gs0")
    for inp, o0, o, o2, rev, run in ((a, ref[0], got[0], got2[0], False, runs[0]), (b, ref[1], got[1], got2[1], True, runs[1])):
        if not run:
            assert o[0] is None
            continue
        for i in range(nout):
            if names[i] == "gu":     # partial sums of a cut sequence are added in another order: bf16 rounding of the sum
                assert oracle.rel_err(o[i], o0[i].double().cpu()) <= 1e-2, names[i]
            else:
                assert torch.equal(o[i], o0[i]), names[i]
            assert torch.equal(o[i], o2[i]), names[i] + " (second call)"


def test_balanced_grid_against_oracle():
    """The bench shape at a shorter T against the fp64 oracle (the oracle's backward is O(T N^2) per sequence on the host)."""
    B, T, H = 20, 200, 8
    (a, b), _, got, _, _ = _run(B, T, H, 3, True, seed=71)
    for inp, o, rev in ((a, got[0], False), (b, got[1], True)):
        ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], s0=inp.get("s0"), gsT=inp.get("gsT"),
                         rev=rev, prec="f64")
        for x, r_ in zip(o[:6], ref[:6]):
            assert oracle.rel_err(x, r_) <= TOL[torch.bfloat16]


def test_balanced_grid_under_cuda_graph_replay():
    """A captured launch carries ONE epoch for all its replays: the consumer CTAs put every hand-over flag back to 0, so a
    replay waits for its own producers instead of reading the previous replay's records.  Inputs change between replays."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 16, 300, 8
    a = _inputs(B, T, H, torch.bfloat16, seed=81)
    b = _inputs(B, T, H, torch.bfloat16, seed=82, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    ck = ops.checkpoint_buffer(ga["r"], 2)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):                    # warm-up outside the capture (lazy driver initialisation)
        pk.bi_forward_raw(f, r_, ckpt=ck)
        pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
    torch.cuda.current_stream().wait_stream(side)
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        pk.bi_forward_raw(f, r_, ckpt=ck)
        of, ob = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
    for scale in (1.0, 0.5, 2.0):
        ga["gy"].copy_((a["gy"].float() * scale).to(torch.bfloat16))
        gb["gy"].copy_((b["gy"].float() * scale).to(torch.bfloat16))
        g.replay()
        torch.cuda.synchronize()
        got = [t.clone() for t in of[:5] + ob[:5]]
        pk.bi_forward_raw(f, r_, ckpt=ck)
        ef, eb = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
        for x, y in zip(got, ef[:5] + eb[:5]):
            assert torch.equal(x, y)
