"""Generates tests/golden/refconv_*.npz by running the reference's OWN ConvolutionModule, imported unmodified from
/root/reference/wenet/transformer/convolution.py in the build container (CPU, fp32): the configuration every RWKV config
uses (kernel 31, layer_norm, swish, non-causal: conf/rwkv/*.yaml:14-18) at configs[0]'s and configs[1]'s widths, with a
padding mask, plus the causal / cache branch and kernel 15.

    python tests/golden/make_refconv_golden.py

Parameters and inputs are regenerated from seeds by the tests (the mirror in paper_accurate_fast_cheap_b200/convolution.py
creates the same torch modules in the same order; this script asserts that)."""
import importlib.util
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
REF = "/root/reference/wenet/transformer/convolution.py"

PARAM_SEED, X_SEED = 31337, 7
# name, C, K, causal, B, T, lengths (None: no mask), with_cache, stored time stride
CASES = [
    ("cfg1_c256_k31", 256, 31, False, 2, 250, [250, 200], False, 1),
    ("cfg2_c512_k31", 512, 31, False, 3, 750, [750, 600, 33], False, 3),
    ("c128_k15", 128, 15, False, 2, 40, None, False, 1),
    ("c256_k31_causal", 256, 31, True, 2, 70, None, False, 1),
    ("c256_k15_causal_cache", 256, 15, True, 2, 50, [50, 45], True, 1),
    ("c1024_k31_short", 1024, 31, False, 1, 9, None, False, 1),
]


def inputs(C, B, T, lengths, with_cache, K):
    g = torch.Generator().manual_seed(X_SEED)
    x = torch.randn(B, T, C, generator=g)
    mask = None
    if lengths is not None:
        mask = (torch.arange(T).view(1, 1, T) < torch.tensor(lengths).view(B, 1, 1))
    cache = torch.randn(B, C, K - 1, generator=g) if with_cache else None
    return x, mask, cache


def randomize(m):
    """Default init leaves LayerNorm at (1, 0): give every parameter a seeded non-trivial value (same order in both modules)."""
    g = torch.Generator().manual_seed(PARAM_SEED + 1)
    with torch.no_grad():
        for p in m.parameters():
            p.add_(0.1 * torch.randn(p.shape, generator=g))


def main():
    spec = importlib.util.spec_from_file_location("ref_convolution", REF)
    ref = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(ref)
    from paper_accurate_fast_cheap_b200.convolution import ConvolutionModule as Mine
    for name, C, K, causal, B, T, lengths, with_cache, stride in CASES:
        torch.manual_seed(PARAM_SEED)
        m = ref.ConvolutionModule(C, K, torch.nn.SiLU(), "layer_norm", causal, True)
        torch.manual_seed(PARAM_SEED)
        mm = Mine(C, K, torch.nn.SiLU(), "layer_norm", causal, True)
        randomize(m); randomize(mm)
        sd, sd2 = m.state_dict(), mm.state_dict()
        assert list(sd) == list(sd2) and all(torch.equal(sd[k], sd2[k]) for k in sd), name
        x, mask, cache = inputs(C, B, T, lengths, with_cache, K)
        m.eval()
        with torch.no_grad():
            kw = {}
            if mask is not None:
                kw["mask_pad"] = mask
            if cache is not None:
                kw["cache"] = cache
            out, new_cache = m(x.clone(), **kw)
            out2, nc2 = mm(x.clone(), **kw)             # the mirror's eager branch on CPU: same ops, same bits
        assert torch.equal(out, out2) and torch.equal(new_cache, nc2), name
        np.savez_compressed(os.path.join(HERE, f"refconv_{name}.npz"), out=out[:, ::stride].numpy(), new_cache=new_cache.numpy(),
                            meta=np.array([C, K, int(causal), B, T, int(with_cache), stride], dtype=np.int64),
                            lengths=np.array(lengths if lengths is not None else [], dtype=np.int64),
                            param_checksum=np.array(float(sum(v.double().abs().sum() for v in sd.values()))))
        print("wrote", name, tuple(out.shape))


if __name__ == "__main__":
    main()
