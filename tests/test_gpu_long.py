"""GPU parity at the full sequence lengths of BASELINE.json configs[2] (long-form, T = 30 000, bf16, d_model 512) and
configs[4] (stress, T = 90 000, d_model 1024 / 16 heads, fp32): direct comparison with the fp64 oracle on a bounded
batch (the oracle is OpenMP C and finishes these in seconds) plus the size-independent property the domain offers --
a run split into several op calls with the state carried between them equals the single run."""
import pytest
import torch

import oracle
from test_gpu_parity import TOL, _cuda, _inputs

pytestmark = pytest.mark.gpu


def _pk():
    import paper_accurate_fast_cheap_b200 as pk
    return pk


@pytest.mark.parametrize("rev", [False, True])
def test_cfg3_longform_bf16_full_length(rev):
    """configs[2]: T = 30 000 subsampled frames, C = 512, 8 heads, bf16; B = 2 of the 8 utterances (one GPU's share
    at 4 GPUs).  y and S_T against the fp64 oracle; tolerance 2e-2 (bf16 I/O)."""
    pk = _pk()
    B, T, H = 2, 30000, 8
    inp = _inputs(B, T, H, torch.bfloat16, seed=31)
    g = _cuda(inp)
    y, sT = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], reverse=rev, want_state=True)
    y_ref, sT_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], rev=rev, want_state=True, prec="f64")
    assert oracle.rel_err(y, y_ref) <= TOL[torch.bfloat16]
    assert oracle.rel_err(sT, sT_ref) <= TOL[torch.bfloat16]


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_cfg3_state_carry_across_calls(dtype):
    """Chunked state carry: three op calls over [0, 9 000), [9 000, 20 037), [20 037, 30 000) with S_T -> S_0 between
    them reproduce the single call.  fp32: to 1e-5; bf16: pieces that do not start on a 64-step boundary regroup the
    tensor-core chunks, so agreement is to the bf16 tolerance, and the carried state to 2e-2."""
    pk = _pk()
    B, T, H = 1, 30000, 8
    inp = _inputs(B, T, H, dtype, seed=32)
    g = _cuda(inp)
    y_all, s_all = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], want_state=True)
    cuts = [0, 9000, 20037, T]
    s, ys = None, []
    for a, b in zip(cuts[:-1], cuts[1:]):
        sl = lambda t: t[:, a:b].contiguous()
        y, s = pk.wkv6_forward_raw(sl(g["r"]), sl(g["k"]), sl(g["v"]), sl(g["w"]), g["u"], s0=s, want_state=True)
        ys.append(y)
    y_cat = torch.cat(ys, dim=1)
    tol = TOL[dtype]
    assert oracle.rel_err(y_cat.cpu(), y_all.cpu().double()) <= tol
    assert oracle.rel_err(s.cpu(), s_all.cpu().double()) <= tol


def test_cfg5_stress_fp32_full_length():
    """configs[4]: one 1-hour utterance, T = 90 000, d_model 1024 = 16 heads x 64, fp32 I/O: full-T forward (both
    directions in one launch) against the fp64 oracle at the fp32 tolerance 1e-5."""
    pk = _pk()
    B, T, H = 1, 90000, 16
    a = _inputs(B, T, H, torch.float32, seed=51)
    b = _inputs(B, T, H, torch.float32, seed=52, layer=9)
    ga, gb = _cuda(a), _cuda(b)
    y_f, y_b, s_f, s_b = pk.bi_forward_raw(tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu"), want_state=True)
    ya, sa = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], want_state=True, prec="f64")
    yb, sb = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], rev=True, want_state=True, prec="f64")
    assert oracle.rel_err(y_f, ya) <= TOL[torch.float32] and oracle.rel_err(y_b, yb) <= TOL[torch.float32]
    assert oracle.rel_err(s_f, sa) <= TOL[torch.float32] and oracle.rel_err(s_b, sb) <= TOL[torch.float32]


def test_cfg5_stress_fp32_gradients_prefix():
    """configs[4] gradient check on a T = 8 192 prefix: every gradient against the fp64 oracle at 1e-5."""
    pk = _pk()
    B, T, H = 1, 8192, 16
    inp = _inputs(B, T, H, torch.float32, seed=53)
    g = _cuda(inp)
    out = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"])
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    for name, x, y in zip(("gr", "gk", "gv", "gw", "gu"), out[:5], ref[:5]):
        assert oracle.rel_err(x, y) <= TOL[torch.float32], name


def test_cfg4_training_shape_bf16_gradients():
    """configs[3] per-GPU share: B = 8 x T = 750, C = 512, 8 heads, bf16, both directions, forward with checkpoints +
    ONE backward launch; every gradient of both directions against the fp64 oracle at 2e-2."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 8, 750, 8
    a = _inputs(B, T, H, torch.bfloat16, seed=41)
    b = _inputs(B, T, H, torch.bfloat16, seed=42, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    ck = ops.checkpoint_buffer(ga["r"], 2)
    pk.bi_forward_raw(f, r_, ckpt=ck)
    of, ob = pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
    for inp, o, rev in ((a, of, False), (b, ob, True)):
        ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], rev=rev, prec="f64")
        for name, x, y in zip(("gr", "gk", "gv", "gw", "gu"), o[:5], ref[:5]):
            assert oracle.rel_err(x, y) <= TOL[torch.bfloat16], (name, rev)


@pytest.mark.parametrize("with_state", [False, True])
def test_cfg3_time_segmented_forward(with_state):
    """configs[2] at 8 GPUs: ONE utterance per GPU (B = 1, 8 heads, both directions = 16 sequences of T = 30 000).  With
    a scratch buffer (ops._scratch, per call) the forward runs time-segmented (pass 1 / segment scan / pass 2: three
    launches); y, S_T and the chunk checkpoints agree with the one-CTA-per-sequence run and with the fp64 oracle."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 1, 30000, 8
    a = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=33)
    b = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=34, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    kw = dict(s0_f=ga.get("s0"), s0_b=gb.get("s0"), want_state=True)
    ops.set_segmentation(False)
    try:
        n0 = pk.launch_count()
        y_f0, y_b0, s_f0, s_b0 = pk.bi_forward_raw(f, r_, **kw)
        assert pk.launch_count() - n0 == 1
    finally:
        ops.set_segmentation(True)
    n0 = pk.launch_count()
    y_f, y_b, s_f, s_b = pk.bi_forward_raw(f, r_, **kw)
    assert pk.launch_count() - n0 == 3, "expected pass 1 + segment scan + pass 2"
    for x, x0 in ((y_f, y_f0), (y_b, y_b0), (s_f, s_f0), (s_b, s_b0)):
        assert oracle.rel_err(x.cpu(), x0.cpu().double()) <= 1e-2      # same tolerance class: bf16 operands, regrouped sums
    ya, sa = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], a.get("s0"), want_state=True, prec="f64")
    yb, sb = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], b.get("s0"), rev=True, want_state=True, prec="f64")
    tol = TOL[torch.bfloat16]
    assert oracle.rel_err(y_f, ya) <= tol and oracle.rel_err(y_b, yb) <= tol
    assert oracle.rel_err(s_f, sa) <= tol and oracle.rel_err(s_b, sb) <= tol


def test_time_segmented_ragged_and_speed():
    """Ragged T (not a multiple of 64, segments of unequal length), single direction; and the point of it: the segmented
    run must be faster than one CTA per sequence on this shape."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 1, 20011, 4
    inp = _inputs(B, T, H, torch.bfloat16, seed=35)
    g = _cuda(inp)
    args = (g["r"], g["k"], g["v"], g["w"], g["u"])

    def timed():
        pk.wkv6_forward_raw(*args, reverse=True, want_state=True)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        out = pk.wkv6_forward_raw(*args, reverse=True, want_state=True)
        e1.record()
        torch.cuda.synchronize()
        return out, e0.elapsed_time(e1)

    ops.set_segmentation(False)
    try:
        (y0, s0_), t_plain = timed()
    finally:
        ops.set_segmentation(True)
    (y1, s1_), t_seg = timed()
    y_ref, s_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], rev=True, want_state=True, prec="f64")
    assert oracle.rel_err(y1, y_ref) <= TOL[torch.bfloat16] and oracle.rel_err(s1_, s_ref) <= TOL[torch.bfloat16]
    assert oracle.rel_err(y0, y_ref) <= TOL[torch.bfloat16]
    print(f"T={T} 4 sequences: one CTA per sequence {t_plain:.3f} ms, time-segmented {t_seg:.3f} ms")
    assert t_seg < 0.5 * t_plain


@pytest.mark.parametrize("with_state", [False, True])
def test_time_segmented_fp32_scan(with_state):
    """The exact fp32 scan takes the same time-segmented route (configs[4]: one 1-hour utterance per GPU = 32 sequences
    on 148 SMs): three launches, y and S_T within the fp32 tolerance of the fp64 oracle and of the unsegmented run,
    ragged T, both directions, with and without an initial state; and it is faster."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 1, 20011, 4
    a = _inputs(B, T, H, torch.float32, with_state=with_state, seed=61)
    b = _inputs(B, T, H, torch.float32, with_state=with_state, seed=62, layer=5)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    kw = dict(s0_f=ga.get("s0"), s0_b=gb.get("s0"), want_state=True)

    def timed():
        pk.bi_forward_raw(f, r_, **kw)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        n0 = pk.launch_count()
        e0.record()
        out = pk.bi_forward_raw(f, r_, **kw)
        e1.record()
        torch.cuda.synchronize()
        return out, e0.elapsed_time(e1), pk.launch_count() - n0

    ops.set_segmentation(False)
    try:
        plain, t_plain, n_plain = timed()
    finally:
        ops.set_segmentation(True)
    seg, t_seg, n_seg = timed()
    assert n_plain == 1 and n_seg == 3
    ya, sa = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], a.get("s0"), want_state=True, prec="f64")
    yb, sb = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], b.get("s0"), rev=True, want_state=True, prec="f64")
    tol = TOL[torch.float32]
    for got in (plain, seg):
        assert oracle.rel_err(got[0], ya) <= tol and oracle.rel_err(got[1], yb) <= tol
        assert oracle.rel_err(got[2], sa) <= tol and oracle.rel_err(got[3], sb) <= tol
    print(f"fp32 scan, 8 sequences of T={T}: one CTA per sequence {t_plain:.3f} ms, time-segmented {t_seg:.3f} ms")
    assert t_seg < 0.6 * t_plain


def test_time_segmented_two_segments_many_sequences():
    """configs[2] at one GPU (8 utterances = 128 sequences): 2 segments per sequence (pass 1 is state-only, so two passes
    over half the length still beat one CTA per sequence on 128 of 296 slots).  Parity as for every segmented launch."""
    pk = _pk()
    B, T, H = 8, 4200, 8
    a = _inputs(B, T, H, torch.bfloat16, with_state=True, seed=71)
    b = _inputs(B, T, H, torch.bfloat16, seed=72, layer=3)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    n0 = pk.launch_count()
    y_f, y_b, s_f, s_b = pk.bi_forward_raw(f, r_, s0_f=ga["s0"], want_state=True)
    assert pk.launch_count() - n0 == 3
    ya, sa = oracle.fwd(a["r"], a["k"], a["v"], a["w"], a["u"], a["s0"], want_state=True, prec="f64")
    yb, sb = oracle.fwd(b["r"], b["k"], b["v"], b["w"], b["u"], rev=True, want_state=True, prec="f64")
    tol = TOL[torch.bfloat16]
    assert oracle.rel_err(y_f, ya) <= tol and oracle.rel_err(y_b, yb) <= tol
    assert oracle.rel_err(s_f, sa) <= tol and oracle.rel_err(s_b, sb) <= tol


@pytest.mark.parametrize("dtype", [torch.bfloat16, torch.float32])
def test_time_segmented_few_heads_long_T(dtype):
    """B = 1, 4 heads, ONE direction, T = 40 000 (the configs[0] model size on a long recording): 4 sequences would ask for
    70 segments; the planner clamps to the 40 the segment-scan kernel handles (this shape used to fail with
    'invalid argument' after pass 1 had been launched).  Three launches, parity with the fp64 oracle."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import _lib
    B, T, H = 1, 40000, 4
    inp = _inputs(B, T, H, dtype, seed=81)
    g = _cuda(inp)
    nb = _lib.lib().wkv6_b200_forward_scratch_bytes(0 if dtype == torch.bfloat16 else 1, B, T, H * 64, H, 1)
    assert 0 < nb <= 4 * (40 * (64 * 64 + 64) + 625 * 64) * 4          # <= 40 segments (+ the per-chunk decay prefixes)
    n0 = pk.launch_count()
    y, sT = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], want_state=True)
    torch.cuda.synchronize()
    assert pk.launch_count() - n0 == 3
    y_ref, s_ref = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], want_state=True, prec="f64")
    assert oracle.rel_err(y, y_ref) <= TOL[dtype] and oracle.rel_err(sT, s_ref) <= TOL[dtype]


def test_scratch_is_per_call_two_streams():
    """ABI v3: the segmentation scratch is an argument of the call, not library state.  Two segmented forwards on two
    streams at the same time (the reference's forward_parallel pattern, rwkv_wrapper_bidirectional.py:66-116) use two
    buffers and both match the serial results; a too-small or NULL scratch runs unsegmented (1 launch), never fails."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import _lib
    L = _lib.lib()
    B, T, H = 1, 20011, 4
    a, b = _cuda(_inputs(B, T, H, torch.bfloat16, seed=91)), _cuda(_inputs(B, T, H, torch.bfloat16, seed=92, layer=4))
    ya, _ = pk.wkv6_forward_raw(a["r"], a["k"], a["v"], a["w"], a["u"])
    yb, _ = pk.wkv6_forward_raw(b["r"], b["k"], b["v"], b["w"], b["u"], reverse=True)
    torch.cuda.synchronize()
    s1, s2 = torch.cuda.Stream(), torch.cuda.Stream()
    outs = []
    for _ in range(4):                      # several rounds so that the two triples really interleave
        with torch.cuda.stream(s1):
            y1, _ = pk.wkv6_forward_raw(a["r"], a["k"], a["v"], a["w"], a["u"])
        with torch.cuda.stream(s2):
            y2, _ = pk.wkv6_forward_raw(b["r"], b["k"], b["v"], b["w"], b["u"], reverse=True)
        outs.append((y1, y2))
    torch.cuda.synchronize()
    for y1, y2 in outs:
        assert torch.equal(y1, ya) and torch.equal(y2, yb)
    # NULL / undersized scratch: one launch, same tolerance class
    p = lambda t: t.data_ptr()
    y = torch.empty_like(a["r"])
    small = torch.empty(1024, device="cuda", dtype=torch.uint8)
    for sc, nb in ((None, 0), (small.data_ptr(), small.numel())):
        n0 = pk.launch_count()
        _lib.check(L.wkv6_b200_forward(0, B, T, H * 64, H, p(a["r"]), p(a["k"]), p(a["v"]), p(a["w"]), p(a["u"]), None, p(y),
                                       None, None, 0, sc, nb, torch.cuda.current_stream().cuda_stream))
        torch.cuda.synchronize()
        assert pk.launch_count() - n0 == 1
        assert oracle.rel_err(y.cpu(), ya.cpu().double()) <= 1e-2
    # misaligned scratch is rejected before any launch
    big = torch.empty((8 << 20) + 16, device="cuda", dtype=torch.uint8)
    rc = L.wkv6_b200_forward(0, B, T, H * 64, H, p(a["r"]), p(a["k"]), p(a["v"]), p(a["w"]), p(a["u"]), None, p(y), None, None, 0,
                             big.data_ptr() + 4, 8 << 20, torch.cuda.current_stream().cuda_stream)
    assert rc == -1


@pytest.mark.skipif(not torch.cuda.is_available() or torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_two_devices_one_process():
    """the dynamic-shared-memory opt-in is a per-device function attribute; one process driving cuda:0 and then
    cuda:1 must get it on both (bf16 forward + backward on each device against the oracle)."""
    pk = _pk()
    inp = _inputs(2, 200, 4, torch.bfloat16, seed=95)
    y_ref, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], prec="f64")
    ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    for d in (0, 1):
        g = {n: t.to(f"cuda:{d}") for n, t in inp.items() if isinstance(t, torch.Tensor)}
        y, _ = pk.wkv6_forward_raw(g["r"], g["k"], g["v"], g["w"], g["u"])
        o = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"])
        torch.cuda.synchronize(d)
        assert oracle.rel_err(y, y_ref) <= TOL[torch.bfloat16]
        for x, r_ in zip(o[:5], ref[:5]):
            assert oracle.rel_err(x, r_) <= TOL[torch.bfloat16]


@pytest.mark.parametrize("B,T,H,flags,with_state", [
    (1, 12499, 8, 3, False),     # FT-LFXL long-form fine-tuning shape: 50 000 input frames -> T' = 12 499, one utterance per GPU
    (1, 2113, 2, 3, True),       # ragged T, initial state and gradient into the final state, 4 sequences x 4 segments
    (2, 4100, 4, 2, True),       # one direction dropped: 8 sequences of the right-to-left set
    (1, 2048, 1, 1, False),      # the shortest sequence that is segmented (32 chunks), left-to-right only
])
def test_time_segmented_backward(B, T, H, flags, with_state):
    """Few, long sequences (long-form fine-tuning; the reference itself cannot run T > ctx_len backward, wkv6_cuda.cu:207):
    with the scratch from wkv6_b200_backward_scratch_bytes the checkpointed backward runs time-segmented (segment-local
    reverse states, segment scan, gradients, gu reduction: four launches); every gradient agrees with the
    one-CTA-per-sequence launch and with the fp64 oracle."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import _lib, ops
    a = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=51)
    b = _inputs(B, T, H, torch.bfloat16, with_state=with_state, seed=52, layer=7)
    ga, gb = _cuda(a), _cuda(b)
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    run_f, run_b = bool(flags & 1), bool(flags & 2)
    ndir = int(run_f) + int(run_b)
    assert _lib.lib().wkv6_b200_backward_scratch_bytes(0, B, T, H * 64, H, ndir) > 0
    ck = ops.checkpoint_buffer(ga["r"], 2)
    pk.bi_forward_raw(f, r_, flags=flags, s0_f=ga.get("s0"), s0_b=gb.get("s0"), ckpt=ck)
    kw = dict(flags=flags, s0_f=ga.get("s0"), s0_b=gb.get("s0"), gsT_f=ga.get("gsT"), gsT_b=gb.get("gsT"),
              want_gs0=with_state, ckpt=ck)
    gyf, gyb = (ga["gy"] if run_f else None), (gb["gy"] if run_b else None)
    ops.set_segmentation(False)
    try:
        n0 = pk.launch_count()
        of0, ob0 = pk.bi_backward_raw(f, r_, gyf, gyb, **kw)
        assert pk.launch_count() - n0 == 1
    finally:
        ops.set_segmentation(True)
    n0 = pk.launch_count()
    of, ob = pk.bi_backward_raw(f, r_, gyf, gyb, **kw)
    assert pk.launch_count() - n0 == 4, "expected state pass + segment scan + gradient pass + gu reduction"
    nout = 6 if with_state else 5
    for inp, o, o0, rev, run in ((a, of, of0, False, run_f), (b, ob, ob0, True, run_b)):
        if not run:
            assert o[0] is None
            continue
        for name, x, x0 in zip(("gr", "gk", "gv", "gw", "gu", "gs0"), o[:nout], o0[:nout]):
            assert oracle.rel_err(x, x0.double()) <= 1e-2, (name, rev)      # same operands, fp32 state regrouped per segment
        if B * T * H <= 70000:                                                # the fp64 oracle backward on the smaller cases
            ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], s0=inp.get("s0"),
                             gsT=inp.get("gsT"), rev=rev, prec="f64")
            for name, x, y in zip(("gr", "gk", "gv", "gw", "gu", "gs0"), o[:nout], ref[:nout]):
                assert oracle.rel_err(x, y) <= TOL[torch.bfloat16], (name, rev)


def test_time_segmented_backward_is_faster():
    """The point of it: B = 1 x T = 12 499 x 8 heads, both directions (16 sequences on 148 SMs)."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    B, T, H = 1, 12499, 8
    ga, gb = _cuda(_inputs(B, T, H, torch.bfloat16, seed=61)), _cuda(_inputs(B, T, H, torch.bfloat16, seed=62, layer=7))
    f, r_ = tuple(ga[n] for n in "rkvwu"), tuple(gb[n] for n in "rkvwu")
    ck = ops.checkpoint_buffer(ga["r"], 2)
    pk.bi_forward_raw(f, r_, ckpt=ck)

    def ms(reps=3):
        pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(reps):
            pk.bi_backward_raw(f, r_, ga["gy"], gb["gy"], ckpt=ck)
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    ops.set_segmentation(False)
    try:
        t_plain = ms()
    finally:
        ops.set_segmentation(True)
    t_seg = ms()
    print(f"\nbackward B=1 T=12499 H=8 both directions: one CTA per sequence {t_plain:.3f} ms, time-segmented {t_seg:.3f} ms")
    assert t_seg < 0.5 * t_plain


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs two GPUs in one process")
def test_one_process_two_devices():
    """One host process driving two GPUs (DataParallel-style, a server looping over devices): the dynamic-shared-memory opt-in
    is a per-device function attribute and the tensor-map cache is keyed by pointer, so every kernel must work on cuda:1
    after cuda:0 (a process-wide `static bool` had opted in on the first device only)."""
    pk = _pk()
    from paper_accurate_fast_cheap_b200 import ops
    from paper_accurate_fast_cheap_b200.convolution import convmod_core
    inp = _inputs(2, 300, 4, torch.bfloat16, seed=71)
    ref_y, _ = oracle.fwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], prec="f64")
    ref_g = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], prec="f64")
    for dev in ("cuda:0", "cuda:1", "cuda:0"):
        g = {n: t.to(dev).contiguous() for n, t in inp.items()}
        f = tuple(g[n] for n in "rkvwu")
        ck = ops.checkpoint_buffer(g["r"], 2)
        y_f, y_b, _, _ = pk.bi_forward_raw(f, f, ckpt=ck)
        of, _ = pk.bi_backward_raw(f, f, g["gy"], g["gy"], ckpt=ck)
        assert y_f.device == torch.device(dev)
        assert oracle.rel_err(y_f, ref_y) <= TOL[torch.bfloat16]
        assert oracle.rel_err(of[3], ref_g[3]) <= TOL[torch.bfloat16]
        y32, _ = pk.wkv6_forward_raw(*(t.float() for t in f))                     # exact scan kernels
        assert oracle.rel_err(y32, ref_y) <= 1e-5
        z = torch.randn(2, 100, 512, device=dev)
        out = convmod_core(z, torch.randn(256, 31, device=dev), None, torch.ones(256, device=dev), torch.zeros(256, device=dev), 1e-5, 100, 15)
        assert out.device == torch.device(dev) and bool(torch.isfinite(out).all())
