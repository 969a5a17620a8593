"""Module-level TRAINING parity: input and parameter gradients of the attention-slot modules
(paper_accurate_fast_cheap_b200/wrappers.py) against gradients of the reference's own Python modules
(tests/golden/refpygrad_*.npz, made by make_refpy_grad_golden.py: rwkv_wrapper_bidirectional.py:43-56,
..._direction_dropout.py:59-69, ..._direction_dropout_both.py:53-71, autograd through model.py:136-152), including
the steps where one direction is dropped and its 17 parameters must come back with `grad is None`
(what makes find_unused_parameters=True necessary, wenet/utils/train_utils.py:365-366).

CPU tests check the host logic (autograd wiring of the flip-free prologue, BiWKV6 / Combine packing, the dropout RNG
sequence) with the scan replaced by an oracle-backed autograd.Function -- in the test, never in the product.  GPU tests
run the real kernels (chunked tcgen05 forward with checkpoints + backward for bf16, exact scan for fp32)."""
import glob
import os

import numpy as np
import pytest
import torch

import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "refpygrad_*.npz")))
PARAM_SEED, X_SEED, CALL_SEED, GY_SEED = 4242, 99, 5, 123
BIG, STRIDE = 4096, 7


class _OracleScan(torch.autograd.Function):
    """Test double of the scan with autograd: forward / backward through the fp32 oracle (either direction)."""

    @staticmethod
    def forward(ctx, r, k, v, w, u, rev):
        ctx.rev = bool(rev)
        ctx.save_for_backward(r, k, v, w, u)
        y, _ = oracle.fwd(r, k, v, w, u, rev=rev, prec="f32")
        return y.to(r.dtype)

    @staticmethod
    def backward(ctx, gy):
        r, k, v, w, u = ctx.saved_tensors
        gr, gk, gv, gw, gu, _ = oracle.bwd(r, k, v, w, u, gy.contiguous(), rev=ctx.rev, prec="f32")
        dt = r.dtype
        # like the kernels: [B,C] partials rounded to the I/O dtype, then summed (model.py:151)
        return gr.to(dt), gk.to(dt), gv.to(dt), gw.to(dt), gu.to(dt).sum(0).view(u.shape), None


def _oracle_backend(monkeypatch, w):
    scan = lambda r, k, v, wd, u, rev: _OracleScan.apply(r, k, v, wd, u, rev)

    def bi_wkv6(f, b, run_l2r=True, run_r2l=True):
        return (scan(*f, False) if run_l2r else None), (scan(*b, True) if run_r2l else None)

    def combine(a, b, scale=0.5, round_sum=False, out_float=True):
        s = a + b if round_sum else a.float() + b.float()
        out = s.float() * scale
        return out if out_float else out.to(a.dtype)

    monkeypatch.setattr(w.autograd, "bi_wkv6", bi_wkv6)
    monkeypatch.setattr(w.autograd, "wkv6", lambda r, k, v, wd, u, s0, rev: (scan(r, k, v, wd, u, rev), None))
    monkeypatch.setattr(w.autograd, "RUN_CUDA_RWKV6", lambda r, k, v, wd, u: scan(r, k, v, wd, u, False))
    monkeypatch.setattr(w.autograd, "RUN_CUDA_RWKV6_FP32", lambda r, k, v, wd, u: scan(r, k, v, wd, u, False))
    monkeypatch.setattr(w.autograd, "combine", combine)


def _sample(t):
    t = t.detach().float().reshape(-1).cpu()
    return t[::STRIDE] if t.numel() >= BIG else t


def _run_case(path, monkeypatch, device, tol_out, tol_grad):
    import paper_accurate_fast_cheap_b200.wrappers as w
    classes = {"bi": w.RWKV_TmixWrapper_bidirectional, "drop": w.RWKV_TmixWrapper_bidirectional_direction_dropout,
               "drop_both": w.RWKV_TmixWrapper_bidirectional_direction_dropout_both}
    z = np.load(path)
    C, layer_id, bf16, B, T, ncalls = (int(v) for v in z["meta"])
    for k_ in ("RWKV_ALT_DECODING", "RWKV_BIDIRECTIONAL_LAYERS"):
        monkeypatch.delenv(k_, raising=False)
    torch.manual_seed(PARAM_SEED + layer_id)
    m = classes[str(z["key"])](64, C, 12, "v6", "bi", 2048, bool(bf16), layer_id)
    names = [n for n, _ in m.named_parameters()]
    assert names == [str(n) for n in z["names"]]
    x0 = torch.randn(B, T, C, generator=torch.Generator().manual_seed(X_SEED))
    gy = torch.randn(B, T, C, generator=torch.Generator().manual_seed(GY_SEED)).to(device)
    m = m.to(device).train()
    torch.manual_seed(CALL_SEED)
    outcomes = [int(v) for v in z["outcomes"]]
    checked, worst = set(), {}
    for step in range(ncalls):
        m.zero_grad(set_to_none=True)
        x = x0.clone().to(device).requires_grad_(True)
        out, _ = m(x, x, x)
        (out.float() * gy).sum().backward()
        has_f = [p.grad is not None for n, p in m.named_parameters() if n.startswith("rwkv_wrapper_forward")]
        has_b = [p.grad is not None for n, p in m.named_parameters() if n.startswith("rwkv_wrapper_backward")]
        # a direction either ran (all 17 tensors have gradients) or was dropped (none has)
        assert all(has_f) == any(has_f) and all(has_b) == any(has_b)
        oc = (1 if any(has_f) else 0) | (2 if any(has_b) else 0)
        assert oc == outcomes[step], f"step {step}: directions {oc}, the reference ran {outcomes[step]}"
        if f"s{oc}_step" not in z or int(z[f"s{oc}_step"]) != step:
            continue
        checked.add(oc)
        e = oracle.rel_err(out, torch.from_numpy(z[f"s{oc}_out"]))
        assert e <= tol_out, f"step {step} out: {e}"
        e = oracle.rel_err(x.grad, torch.from_numpy(z[f"s{oc}_gx"]))
        worst["gx"] = max(worst.get("gx", 0.0), e)
        assert e <= tol_grad, f"step {step} x.grad: {e}"
        for i, (n, p) in enumerate(m.named_parameters()):
            if p.grad is None:
                assert f"s{oc}_g{i}" not in z, n
                continue
            e = oracle.rel_err(_sample(p.grad), torch.from_numpy(z[f"s{oc}_g{i}"]))
            worst[n.split(".")[-2] + "." + n.split(".")[-1]] = max(worst.get(n, 0.0), e)
            assert e <= tol_grad, f"step {step} grad of {n}: {e}"
    assert checked == set(outcomes), "every direction outcome of the sequence must have been compared"
    return worst


def test_refpygrad_goldens_present():
    assert len(CASES) >= 5
    seen = set()
    for p in CASES:
        seen |= set(int(v) for v in np.load(p)["outcomes"])
    assert seen == {1, 2, 3}            # both directions, L2R only, R2L only


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[10:-4] for p in CASES])
def test_module_gradients_match_reference_modules_cpu(path, monkeypatch):
    import paper_accurate_fast_cheap_b200.wrappers as w
    _oracle_backend(monkeypatch, w)
    bf16 = bool(int(np.load(path)["meta"][2]))
    # same torch CPU ops; freedom: the flip-free token shift / summation order over time, and (bf16) the rounding of
    # intermediate sums
    _run_case(path, monkeypatch, "cpu", 1e-2 if bf16 else 2e-6, 4e-2 if bf16 else 5e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[10:-4] for p in CASES])
def test_module_gradients_match_reference_modules_gpu(path, monkeypatch):
    bf16 = bool(int(np.load(path)["meta"][2]))
    worst = _run_case(path, monkeypatch, "cuda:0", 3e-2 if bf16 else 1e-4, 5e-2 if bf16 else 3e-4)
    print(os.path.basename(path), "worst relative errors:", {k: f"{v:.2e}" for k, v in sorted(worst.items())})


@pytest.mark.gpu
def test_graphed_training_matches_eager_with_direction_dropout():
    """GraphedTraining: forward and backward of a direction-dropout module as CUDA graphs, one pair per direction outcome,
    drawn on the host exactly like the eager module.  Over a seeded sequence that takes all three outcomes, outputs and every
    input / parameter gradient equal the eager step's bit for bit, and a dropped direction's parameters get no gradient."""
    import paper_accurate_fast_cheap_b200 as pk
    torch.manual_seed(21)
    m = pk.RWKV_TmixWrapper_bidirectional_direction_dropout_both(64, 128, 12, "v6", "bi", 2048, True, 3).cuda().train()
    with torch.no_grad():
        for blk in m._blocks():
            blk.time_maa_rkvw_w1.normal_(0, 0.05)
            blk.time_decay_w1.normal_(0, 0.05)
    xs = [torch.randn(2, 130, 128, device="cuda") for _ in range(2)]
    gy = torch.randn(2, 130, 128, device="cuda")
    slot = pk.GraphedTraining(m, xs[0])

    def run(step_fn, seed):
        torch.manual_seed(seed)
        res = []
        for i in range(10):
            m.zero_grad(set_to_none=True)
            x = xs[i & 1].clone().requires_grad_(True)
            out = step_fn(x)
            (out.float() * gy).sum().backward()
            res.append((out.detach().clone(), x.grad.clone(), [None if p.grad is None else p.grad.clone() for p in m.parameters()]))
        return res

    eager = run(lambda x: m(x, x, x)[0], 5)
    graphed = run(slot, 5)
    outcomes = set()
    for (o1, gx1, gp1), (o2, gx2, gp2) in zip(eager, graphed):
        assert torch.equal(o1, o2) and torch.equal(gx1, gx2)
        for (name, _), a, b in zip(m.named_parameters(), gp1, gp2):
            assert (a is None) == (b is None)
            if a is None:
                continue
            if "time_maa" in name:
                # the token-shift mix vectors' gradients are column sums accumulated with fp32 atomics across CTAs
                # (csrc/wkv6_tmix.cu): the order of the additions is not fixed, so two runs may differ in the last bf16 bit
                assert (a.float() - b.float()).abs().max() <= 2.0 ** -7 * a.float().abs().max(), name
            else:
                assert torch.equal(a, b), name
        n_f = sum(g is not None for (n, _), g in zip(m.named_parameters(), gp1) if n.startswith("rwkv_wrapper_forward"))
        n_b = sum(g is not None for (n, _), g in zip(m.named_parameters(), gp1) if n.startswith("rwkv_wrapper_backward"))
        outcomes.add((n_f > 0, n_b > 0))
    assert outcomes == {(True, True), (True, False), (False, True)} and len(slot._graphs) == 3
