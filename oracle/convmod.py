"""oracle/convmod.py -- CPU restatement (numpy, fp64 or fp32) of the Conformer ConvolutionModule forward,
/root/reference/wenet/transformer/convolution.py:89-144, in the [B, T, C] layout.  TEST INFRASTRUCTURE, NOT PRODUCT CODE:
only tests/, __graft_entry__.smoke() and bench.py's CPU legs may import it.

Pinned by tests/golden/refconv_*.npz, outputs of the reference's own module imported from /root/reference
(tests/golden/make_refconv_golden.py)."""
import numpy as np


def sigmoid(x):
    return 1.0 / (1.0 + np.exp(-x))


def convmod_core(z, dw_weight, dw_bias, ln_weight, ln_bias, eps, t_out, left):
    """GLU (:123), depthwise Conv1d (:126; cross-correlation, zero padding), LayerNorm over channels (:127-131),
    swish (:129).  z [B, Tin, 2C]; dw_weight [C, K]; returns [B, t_out, C]."""
    B, Tin, C2 = z.shape
    C = C2 // 2
    K = dw_weight.shape[1]
    g = z[..., :C] * sigmoid(z[..., C:])
    gp = np.zeros((B, left + Tin + K + t_out, C), dtype=z.dtype)
    gp[:, left:left + Tin] = g
    d = np.zeros((B, t_out, C), dtype=z.dtype)
    for k in range(K):
        d += gp[:, k:k + t_out] * dw_weight[:, k][None, None, :]
    if dw_bias is not None:
        d += dw_bias[None, None, :]
    mean = d.mean(-1, keepdims=True)
    var = ((d - mean) ** 2).mean(-1, keepdims=True)
    y = (d - mean) / np.sqrt(var + eps) * ln_weight + ln_bias
    return y * sigmoid(y)


def conv_module_forward(x, params, kernel_size, causal=False, mask_pad=None, cache=None, dtype=np.float64):
    """Whole module (:89-144) for norm='layer_norm', activation swish.  x [B, T, C]; mask_pad [B, 1, T] bool or None;
    cache [B, C, lorder] or None.  params: dict with the reference's state-dict names.  Returns (out [B,T,C], new_cache)."""
    f = lambda a: None if a is None else np.asarray(a, dtype=dtype)
    x = f(x)
    B, T, C = x.shape
    if mask_pad is not None:
        x = x * np.asarray(mask_pad, dtype=bool).transpose(0, 2, 1)          # :107-108
    lorder = kernel_size - 1 if causal else 0
    if lorder > 0:                                                            # :110-118
        pad = np.zeros((B, lorder, C), dtype=dtype) if cache is None else f(cache).transpose(0, 2, 1)
        x = np.concatenate([pad, x], axis=1)
        new_cache = x[:, -lorder:].transpose(0, 2, 1)
        left = 0
    else:
        new_cache = np.zeros((0, 0, 0), dtype=dtype)
        left = (kernel_size - 1) // 2
    w1 = f(params["pointwise_conv1.weight"])[:, :, 0]
    z = x @ w1.T + (f(params["pointwise_conv1.bias"]) if "pointwise_conv1.bias" in params else 0.0)    # :122
    y = convmod_core(z, f(params["depthwise_conv.weight"])[:, 0, :], f(params.get("depthwise_conv.bias")),
                     f(params["norm.weight"]), f(params["norm.bias"]), 1e-5, T, left)
    w2 = f(params["pointwise_conv2.weight"])[:, :, 0]
    out = y @ w2.T + (f(params["pointwise_conv2.bias"]) if "pointwise_conv2.bias" in params else 0.0)  # :138
    if mask_pad is not None:
        out = out * np.asarray(mask_pad, dtype=bool).transpose(0, 2, 1)       # :140-141
    return out, new_cache
