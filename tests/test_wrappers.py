"""Host mirror of the reference's attention-slot modules (paper_accurate_fast_cheap_b200/wrappers.py) against golden
outputs of the reference's own Python modules (tests/golden/refpy_*.npz, made by make_refpy_golden.py).

CPU tests check the HOST logic only (parameter init / names, flip-free prologue, direction-dropout RNG sequence,
env switches, combine semantics): the scan is replaced by the oracle through monkeypatching -- in the test, never
in the product.  GPU tests run the real kernels through the same modules."""
import glob
import os

import numpy as np
import pytest
import torch

import oracle

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "refpy_*.npz")))
PARAM_SEED, X_SEED, CALL_SEED = 4242, 99, 5


def _classes():
    import paper_accurate_fast_cheap_b200.wrappers as w
    return w, {"uni": w.RWKV_TmixWrapper, "bi": w.RWKV_TmixWrapper_bidirectional,
               "drop": w.RWKV_TmixWrapper_bidirectional_direction_dropout,
               "drop_both": w.RWKV_TmixWrapper_bidirectional_direction_dropout_both}


def _oracle_backend(monkeypatch, w):
    def scan(r, k, v, wd, u, rev):
        y, _ = oracle.fwd(r, k, v, wd, u, rev=rev, prec="f32")
        return y.to(r.dtype)

    def bi_wkv6(f, b, run_l2r=True, run_r2l=True):
        return (scan(*f, False) if run_l2r else None), (scan(*b, True) if run_r2l else None)

    def combine(a, b, scale=0.5, round_sum=False, out_float=True):
        s = a + b if round_sum else a.float() + b.float()
        return s.float() * scale

    monkeypatch.setattr(w.autograd, "bi_wkv6", bi_wkv6)
    monkeypatch.setattr(w.autograd, "wkv6", lambda r, k, v, wd, u, s0, rev: (scan(r, k, v, wd, u, rev), None))
    monkeypatch.setattr(w.autograd, "RUN_CUDA_RWKV6", lambda r, k, v, wd, u: scan(r, k, v, wd, u, False))
    monkeypatch.setattr(w.autograd, "RUN_CUDA_RWKV6_FP32", lambda r, k, v, wd, u: scan(r, k, v, wd, u, False))
    monkeypatch.setattr(w.autograd, "combine", combine)


def _run_case(path, monkeypatch, device="cpu"):
    z = np.load(path)
    C, layer_id, bf16, B, T, ncalls, stride = (int(v) for v in z["meta"])
    w, classes = _classes()
    for k_ in ("RWKV_ALT_DECODING", "RWKV_BIDIRECTIONAL_LAYERS"):
        monkeypatch.delenv(k_, raising=False)
    for kv in str(z["env"]).split(";"):
        if kv:
            monkeypatch.setenv(*kv.split("="))
    torch.manual_seed(PARAM_SEED + layer_id)
    m = classes[str(z["key"])](64, C, 12, "v6", "bi", 2048, bool(bf16), layer_id)
    psum = float(sum(v.double().abs().sum() for v in m.state_dict().values()))
    assert abs(psum - float(z["param_checksum"])) <= 1e-9 * psum, "parameter init drifted from the reference's"
    x = torch.randn(B, T, C, generator=torch.Generator().manual_seed(X_SEED))
    assert abs(float(x.double().abs().sum()) - float(z["x_checksum"])) < 1e-6
    m, xd = m.to(device), x.to(device)
    outs = []
    with torch.no_grad():
        if ncalls == 0:
            m.eval()
            o, cache = m(xd, xd, xd)
            outs.append(o)
        else:
            m.train()
            torch.manual_seed(CALL_SEED)
            for _ in range(ncalls):
                o, cache = m(xd, xd, xd)
                outs.append(o)
    assert cache.numel() == 0                                    # the cache is passed through untouched
    out = torch.stack(outs, 0)[:, :, ::stride].cpu()
    assert out.dtype == torch.float32
    return out, torch.from_numpy(z["out"]), bool(bf16)


def test_refpy_goldens_present():
    assert len(CASES) >= 8


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[6:-4] for p in CASES])
def test_host_mirror_matches_reference_modules_cpu(path, monkeypatch):
    w, _ = _classes()
    _oracle_backend(monkeypatch, w)
    out, gold, bf16 = _run_case(path, monkeypatch)
    # identical torch CPU ops in the same order; the only freedom is the flip-free token shift (exactly equal math)
    assert oracle.rel_err(out, gold) <= (1e-2 if bf16 else 2e-6)


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[6:-4] for p in CASES])
def test_host_mirror_matches_reference_modules_gpu(path, monkeypatch):
    out, gold, bf16 = _run_case(path, monkeypatch, device="cuda:0")
    # fp32 blocks: cuBLAS (TF32 off) vs CPU GEMM ordering + the 1e-5 scan; bf16 blocks: bf16 GEMMs and LayerNorm
    assert oracle.rel_err(out, gold) <= (3e-2 if bf16 else 1e-4)


def test_state_dict_names_match_reference_layout():
    w, classes = _classes()
    m = classes["bi"](64, 128, 12, "v6", "bi", 2048, True, 1)
    names = set(m.state_dict().keys())
    for d in ("rwkv_wrapper_forward", "rwkv_wrapper_backward"):
        for p in ("time_maa_x", "time_maa_r", "time_maa_k", "time_maa_v", "time_maa_w", "time_maa_rkvw_w1",
                  "time_maa_rkvw_w2", "time_decay", "time_decay_w1", "time_decay_w2", "time_faaaa",
                  "receptance.weight", "key.weight", "value.weight", "output.weight", "ln_x.weight", "ln_x.bias"):
            assert f"{d}.tmix_block.{p}" in names
    assert len(names) == 34
    assert m.rwkv_wrapper_forward.tmix_block.receptance.weight.dtype == torch.bfloat16
    assert set(w.WENET_ATTENTION_CLASSES_B200) == {"rwkv_tmix60", "rwkv_tmix60_bidirectional",
                                                   "rwkv_tmix60_bidirectional2", "rwkv_tmix60_dir_layer_drop",
                                                   "rwkv_tmix60_dir_layer_drop_both"}


def test_wrappers_have_no_cpu_scan():
    """Without monkeypatching, the product path must refuse CPU tensors (no fallback)."""
    _, classes = _classes()
    m = classes["bi"](64, 64, 12, "v6", "bi", 2048, False, 1).eval()
    x = torch.randn(1, 8, 64)
    with pytest.raises(Exception):
        m(x, x, x)


@pytest.mark.gpu
def test_graphed_inference_replays_the_modules():
    """The whole attention-slot stack as one CUDA graph: every C-ABI entry point is capturable; replay == eager, bit for bit."""
    import paper_accurate_fast_cheap_b200 as pk
    torch.manual_seed(11)
    mods = [pk.RWKV_TmixWrapper_bidirectional(64, 128, 12, "v6", "bi", 2048, True, l).cuda().eval() for l in range(3)]
    mods.append(pk.RWKV_TmixWrapper(64, 128, 12, "v6", "bi", 2048, True, 5).cuda().eval())
    x1, x2 = torch.randn(2, 70, 128, device="cuda"), torch.randn(2, 70, 128, device="cuda")
    g = pk.GraphedInference(mods, x1)
    n0 = pk.launch_count()
    for x in (x1, x2, x1):
        outs = [o.clone() for o in g(x)]
        with torch.no_grad():
            want = [m(x, x, x)[0] for m in mods]
        for a, b in zip(outs, want):
            assert torch.equal(a, b)
    assert pk.launch_count() > n0                        # the eager reference calls launched; replays do not go through the ABI
