"""Host-side mirror of the reference's Conformer ConvolutionModule:

    wenet/transformer/convolution.py:24-144   ConvolutionModule

Same constructor, parameters / state-dict names (pointwise_conv1, depthwise_conv, norm, pointwise_conv2 are the same
torch modules, created in the same order, so a reference checkpoint loads unchanged and the default initialisation is
RNG-identical), same forward signature `(x, mask_pad, cache) -> (out, new_cache)` and cache semantics.

What is different is HOW the layer_norm + swish configuration (every RWKV config: conf/rwkv/*.yaml:14-18) runs on a
CUDA device without autograd: the input stays in the encoder's [B, T, C] layout -- the two pointwise convolutions are
plain GEMMs on that layout (`F.linear` with the Conv1d weights viewed as matrices) and everything between them (GLU,
depthwise convolution, LayerNorm, swish) is ONE kernel of libwkv6_b200.so (csrc/convmod.cu) instead of seven eager
kernels and two transposes.  Batch-norm, other activations, kernel sizes other than 15 / 31, CPU tensors and training
(autograd on) run the reference's eager sequence of torch ops.
"""
from __future__ import annotations

from typing import Tuple

import torch
import torch.nn.functional as F
from torch import nn

from . import _lib

_FAKE_MASK = torch.ones((0, 0, 0), dtype=torch.bool)
_FAKE_CACHE = torch.zeros((0, 0, 0))


def _is_swish(m: nn.Module) -> bool:
    return isinstance(m, nn.SiLU) or type(m).__name__ == "Swish"


def convmod_core(z: torch.Tensor, dw_weight: torch.Tensor, dw_bias, ln_weight: torch.Tensor, ln_bias: torch.Tensor,
                 eps: float, t_out: int, left: int) -> torch.Tensor:
    """z [B, Tin, 2C] (fp32 or bf16, CUDA, contiguous) -> swish(LayerNorm(depthwise_conv(glu(z)))) [B, t_out, C]."""
    assert z.is_cuda and z.is_contiguous() and z.dim() == 3, "convmod_core: contiguous CUDA tensor [B, Tin, 2C] (no CPU fallback)"
    B, Tin, C2 = z.shape
    C = C2 // 2
    K = dw_weight.shape[-1]
    dt = _lib.DTYPE_BF16 if z.dtype == torch.bfloat16 else _lib.DTYPE_F32
    assert z.dtype in (torch.bfloat16, torch.float32)
    f32 = lambda t: None if t is None else t.detach().reshape(-1).to(device=z.device, dtype=torch.float32).contiguous()
    w, b, lw, lb = f32(dw_weight), f32(dw_bias), f32(ln_weight), f32(ln_bias)
    y = torch.empty((B, t_out, C), device=z.device, dtype=z.dtype)
    p = lambda t: None if t is None else t.data_ptr()
    with torch.cuda.device(z.device):
        _lib.check(_lib.lib().wkv6_b200_convmod_core(dt, B, Tin, t_out, C, K, left, p(z), p(w), p(b), p(lw), p(lb), float(eps),
                                                     p(y), torch.cuda.current_stream(z.device).cuda_stream))
    return y


class ConvolutionModule(nn.Module):
    """ConvolutionModule in Conformer model (convolution.py:24-144)."""

    def __init__(self, channels: int, kernel_size: int = 15, activation: nn.Module = nn.ReLU(), norm: str = "batch_norm",
                 causal: bool = False, bias: bool = True):
        super().__init__()
        self.pointwise_conv1 = nn.Conv1d(channels, 2 * channels, kernel_size=1, stride=1, padding=0, bias=bias)
        if causal:
            padding = 0
            self.lorder = kernel_size - 1
        else:
            assert (kernel_size - 1) % 2 == 0
            padding = (kernel_size - 1) // 2
            self.lorder = 0
        self.depthwise_conv = nn.Conv1d(channels, channels, kernel_size, stride=1, padding=padding, groups=channels, bias=bias)
        assert norm in ["batch_norm", "layer_norm"]
        if norm == "batch_norm":
            self.use_layer_norm = False
            self.norm = nn.BatchNorm1d(channels)
        else:
            self.use_layer_norm = True
            self.norm = nn.LayerNorm(channels)
        self.pointwise_conv2 = nn.Conv1d(channels, channels, kernel_size=1, stride=1, padding=0, bias=bias)
        self.activation = activation
        self.channels, self.kernel_size = channels, kernel_size

    def _fused_ok(self, x: torch.Tensor) -> bool:
        return (x.is_cuda and not torch.is_grad_enabled() and self.use_layer_norm and _is_swish(self.activation)
                and self.kernel_size in (15, 31) and self.channels % 64 == 0 and 64 <= self.channels <= 1024
                and x.dtype in (torch.float32, torch.bfloat16))

    def forward(self, x: torch.Tensor, mask_pad: torch.Tensor = _FAKE_MASK, cache: torch.Tensor = _FAKE_CACHE
                ) -> Tuple[torch.Tensor, torch.Tensor]:
        if not self._fused_ok(x):
            return self._forward_eager(x, mask_pad, cache)
        # ---- [B, T, C] throughout (convolution.py:106-144 without the transposes)
        B, T, C = x.shape
        masked = mask_pad.size(2) > 0
        keep = mask_pad.transpose(1, 2) if masked else None             # [B, T, 1]
        if masked:
            x = x.masked_fill(~keep, 0.0)
        if self.lorder > 0:
            if cache.size(2) == 0:
                x = F.pad(x, (0, 0, self.lorder, 0), "constant", 0.0)
            else:
                assert cache.size(0) == x.size(0) and cache.size(1) == x.size(2)
                x = torch.cat((cache.transpose(1, 2).to(x.dtype), x), dim=1)
            assert x.size(1) > self.lorder
            new_cache = x[:, -self.lorder:].transpose(1, 2)             # (#batch, channels, cache_t), like the reference
            left = 0
        else:
            new_cache = torch.zeros((0, 0, 0), dtype=x.dtype, device=x.device)
            left = (self.kernel_size - 1) // 2
        pw1, pw2 = self.pointwise_conv1, self.pointwise_conv2
        # the reference's pointwise convolutions are cuDNN calls and follow torch.backends.cudnn.allow_tf32 (on by default);
        # the same GEMMs as F.linear follow the matmul switch: keep the reference's precision policy, whatever it is set to
        prev = torch.backends.cuda.matmul.allow_tf32
        torch.backends.cuda.matmul.allow_tf32 = bool(torch.backends.cudnn.allow_tf32) if x.dtype == torch.float32 else prev
        try:
            z = F.linear(x, pw1.weight.view(2 * C, C), pw1.bias).contiguous()               # [B, Tin, 2C]
            y = convmod_core(z, self.depthwise_conv.weight, self.depthwise_conv.bias, self.norm.weight, self.norm.bias,
                             self.norm.eps, T, left)
            out = F.linear(y, pw2.weight.view(C, C), pw2.bias)
        finally:
            torch.backends.cuda.matmul.allow_tf32 = prev
        if masked:
            out = out.masked_fill(~keep, 0.0)
        return out, new_cache

    def _forward_eager(self, x, mask_pad, cache):
        """The reference's sequence of torch ops (training, CPU, batch_norm, other activations)."""
        x = x.transpose(1, 2)
        if mask_pad.size(2) > 0:
            x = x.masked_fill(~mask_pad, 0.0)
        if self.lorder > 0:
            if cache.size(2) == 0:
                x = F.pad(x, (self.lorder, 0), "constant", 0.0)
            else:
                assert cache.size(0) == x.size(0) and cache.size(1) == x.size(1)
                x = torch.cat((cache, x), dim=2)
            assert x.size(2) > self.lorder
            new_cache = x[:, :, -self.lorder:]
        else:
            new_cache = torch.zeros((0, 0, 0), dtype=x.dtype, device=x.device)
        x = self.pointwise_conv1(x)
        x = F.glu(x, dim=1)
        x = self.depthwise_conv(x)
        if self.use_layer_norm:
            x = x.transpose(1, 2)
        x = self.activation(self.norm(x))
        if self.use_layer_norm:
            x = x.transpose(1, 2)
        x = self.pointwise_conv2(x)
        if mask_pad.size(2) > 0:
            x = x.masked_fill(~mask_pad, 0.0)
        return x.transpose(1, 2), new_cache


def install_convolution(convolution_module=None) -> None:
    """Swaps the class in wenet.transformer.convolution (and wenet.transformer.encoder, which imported the name)."""
    if convolution_module is None:
        from wenet.transformer import convolution as convolution_module   # noqa: PLC0415
    convolution_module.ConvolutionModule = ConvolutionModule
    try:
        from wenet.transformer import encoder as enc                       # noqa: PLC0415
        enc.ConvolutionModule = ConvolutionModule
    except Exception:                                                      # noqa: BLE001
        pass
