"""Conformer ConvolutionModule: the oracle restatement and the host mirror against golden outputs of the
reference's own module (tests/golden/refconv_*.npz, make_refconv_golden.py: wenet/transformer/convolution.py:89-144), and
the fused CUDA core (csrc/convmod.cu, through the C ABI) against the oracle on the same inputs.
Tolerance: max|a-b| / max|b| <= 1e-5 for fp32 I/O, <= 2e-2 for bf16 I/O (the project's definition)."""
import glob
import os

import numpy as np
import pytest
import torch

import oracle
from oracle import convmod as oconv

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
CASES = sorted(glob.glob(os.path.join(GOLD, "refconv_*.npz")))
PARAM_SEED, X_SEED = 31337, 7


def _build(path):
    from paper_accurate_fast_cheap_b200.convolution import ConvolutionModule
    z = np.load(path)
    C, K, causal, B, T, with_cache, stride = (int(v) for v in z["meta"])
    torch.manual_seed(PARAM_SEED)
    m = ConvolutionModule(C, K, torch.nn.SiLU(), "layer_norm", bool(causal), True)
    g = torch.Generator().manual_seed(PARAM_SEED + 1)
    with torch.no_grad():
        for p in m.parameters():
            p.add_(0.1 * torch.randn(p.shape, generator=g))
    psum = float(sum(v.double().abs().sum() for v in m.state_dict().values()))
    assert abs(psum - float(z["param_checksum"])) <= 1e-9 * psum, "parameter init drifted from the reference's"
    g = torch.Generator().manual_seed(X_SEED)
    x = torch.randn(B, T, C, generator=g)
    lengths = [int(v) for v in z["lengths"]]
    mask = (torch.arange(T).view(1, 1, T) < torch.tensor(lengths).view(B, 1, 1)) if lengths else None
    cache = torch.randn(B, C, K - 1, generator=g) if with_cache else None
    return m.eval(), x, mask, cache, z, (C, K, bool(causal), stride)


def test_refconv_goldens_present():
    assert len(CASES) >= 6


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[8:-4] for p in CASES])
def test_oracle_matches_reference_module(path):
    m, x, mask, cache, z, (C, K, causal, stride) = _build(path)
    params = {k: v.numpy() for k, v in m.state_dict().items()}
    out, new_cache = oconv.conv_module_forward(x.numpy(), params, K, causal, None if mask is None else mask.numpy(),
                                               None if cache is None else cache.numpy())
    assert oracle.rel_err(torch.from_numpy(out[:, ::stride]), torch.from_numpy(z["out"])) <= 2e-6
    assert new_cache.shape == z["new_cache"].shape
    if new_cache.size:
        assert oracle.rel_err(torch.from_numpy(new_cache), torch.from_numpy(z["new_cache"])) <= 1e-6


@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[8:-4] for p in CASES])
def test_mirror_eager_branch_is_the_reference_sequence(path):
    m, x, mask, cache, z, (C, K, causal, stride) = _build(path)
    kw = {}
    if mask is not None:
        kw["mask_pad"] = mask
    if cache is not None:
        kw["cache"] = cache
    with torch.no_grad():
        out, new_cache = m(x.clone(), **kw)
    assert torch.equal(out[:, ::stride], torch.from_numpy(z["out"]))
    assert torch.equal(new_cache, torch.from_numpy(z["new_cache"]))
    assert list(m.state_dict()) == ["pointwise_conv1.weight", "pointwise_conv1.bias", "depthwise_conv.weight",
                                    "depthwise_conv.bias", "norm.weight", "norm.bias", "pointwise_conv2.weight",
                                    "pointwise_conv2.bias"]


@pytest.mark.gpu
@pytest.mark.parametrize("path", CASES, ids=[os.path.basename(p)[8:-4] for p in CASES])
def test_mirror_fused_path_matches_reference_module_gpu(path):
    import paper_accurate_fast_cheap_b200 as pk
    m, x, mask, cache, z, (C, K, causal, stride) = _build(path)
    m = m.to("cuda:0")
    kw = {}
    if mask is not None:
        kw["mask_pad"] = mask.to("cuda:0")
    if cache is not None:
        kw["cache"] = cache.to("cuda:0")
    n0 = pk.launch_count()
    with torch.no_grad(), torch.backends.cudnn.flags(allow_tf32=False):     # the pointwise GEMMs follow this switch
        out, new_cache = m(x.to("cuda:0"), **kw)
    assert pk.launch_count() - n0 == 1, "the fused core must have run (one launch)"
    # cuBLAS fp32 GEMMs vs the CPU GEMMs of the golden + the 1e-5 core
    assert oracle.rel_err(out[:, ::stride], torch.from_numpy(z["out"])) <= 1e-4
    assert tuple(new_cache.shape) == z["new_cache"].shape
    if new_cache.numel():
        assert torch.equal(new_cache.cpu(), torch.from_numpy(z["new_cache"]))
    # with autograd on the module takes the eager sequence (no launch of ours) and still agrees
    xg = x.to("cuda:0").requires_grad_(True)
    n0 = pk.launch_count()
    with torch.backends.cudnn.flags(allow_tf32=False):      # cuDNN's Conv1d would otherwise run TF32 (1e-3)
        out2, _ = m(xg, **kw)
    assert pk.launch_count() == n0 and out2.requires_grad
    assert oracle.rel_err(out2[:, ::stride], torch.from_numpy(z["out"])) <= 1e-4


@pytest.mark.gpu
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("B,T,C,K,causal", [
    (2, 250, 256, 31, False),     # configs[0] width
    (16, 750, 512, 31, False),    # configs[1]: the bench shape
    (1, 1, 64, 31, False),        # a single frame, narrowest width
    (3, 7, 128, 15, False),       # shorter than the kernel
    (2, 97, 1024, 31, False),     # configs[4] width, ragged last tile
    (2, 130, 512, 31, True),      # causal: left context in front of the input
    (1, 5000, 512, 31, False),    # long: many tiles per batch row
])
def test_core_matches_oracle(dtype, B, T, C, K, causal):
    from paper_accurate_fast_cheap_b200.convolution import convmod_core
    g = torch.Generator().manual_seed(1234 + T + C)
    t_in = T + K - 1 if causal else T
    z = torch.randn(B, t_in, 2 * C, generator=g).to(dtype)
    w = 0.3 * torch.randn(C, K, generator=g)
    b = 0.1 * torch.randn(C, generator=g)
    lw = 1.0 + 0.2 * torch.randn(C, generator=g)
    lb = 0.2 * torch.randn(C, generator=g)
    left = 0 if causal else (K - 1) // 2
    y = convmod_core(z.to("cuda:0"), w.to("cuda:0"), b.to("cuda:0"), lw.to("cuda:0"), lb.to("cuda:0"), 1e-5, T, left)
    assert y.dtype == dtype and tuple(y.shape) == (B, T, C)
    ref = oconv.convmod_core(z.double().numpy(), w.double().numpy(), b.double().numpy(), lw.double().numpy(),
                             lb.double().numpy(), 1e-5, T, left)
    tol = 1e-5 if dtype == torch.float32 else 2e-2
    assert oracle.rel_err(y, torch.from_numpy(ref)) <= tol


def test_core_rejects_cpu_and_bad_shapes():
    from paper_accurate_fast_cheap_b200 import _lib
    from paper_accurate_fast_cheap_b200.convolution import convmod_core
    L = _lib.lib()
    assert L.wkv6_b200_convmod_core(1, 1, 4, 4, 100, 31, 15, None, None, None, None, None, 1e-5, None, None) == -2
    assert L.wkv6_b200_convmod_core(1, 1, 4, 4, 128, 7, 3, None, None, None, None, None, 1e-5, None, None) == -2
    assert L.wkv6_b200_convmod_core(1, 0, 4, 4, 128, 31, 15, None, None, None, None, None, 1e-5, None, None) == 0
    assert L.wkv6_b200_convmod_core(1, 1, 4, 4, 128, 31, 15, None, None, None, None, None, 1e-5, None, None) == -1
    with pytest.raises(AssertionError):
        convmod_core(torch.zeros(1, 4, 256), torch.zeros(128, 31), None, torch.ones(128), torch.zeros(128), 1e-5, 4, 15)
