"""Streaming / state carry through the attention slot: RWKV_TmixWrapper with a non-empty `cache` carries the
WKV state and the last input frame from piece to piece; the concatenated outputs of the pieces equal one call over the
whole utterance (the domain's size-independent property), and the reference's behaviour -- an empty
cache is passed through untouched (rwkv_wrapper.py:81) -- is unchanged."""
import pytest
import torch

import oracle

CUTS = [0, 100, 173, 300]


def _wrapper(bf16, device):
    import paper_accurate_fast_cheap_b200.wrappers as w
    torch.manual_seed(11)
    m = w.RWKV_TmixWrapper(64, 256, 12, "v6", "l2r", 2048, bf16, 4)
    with torch.no_grad():                                   # the LoRA / decay-LoRA factors start at zero: make them matter
        m.tmix_block.time_maa_rkvw_w1.normal_(0, 0.05)
        m.tmix_block.time_decay_w1.normal_(0, 0.05)
    return w, m.to(device).eval()


def _stream_vs_whole(m, x):
    with torch.no_grad():
        whole, c0 = m(x, x, x)
        assert c0.numel() == 0
        cache = m.init_cache(x.shape[0], x.device)
        outs = []
        for a, b in zip(CUTS[:-1], CUTS[1:]):
            o, cache = m(x[:, a:b].contiguous(), None, None, cache=cache)
            outs.append(o)
        first, _ = m(x[:, :CUTS[1]].contiguous())
    assert cache.shape == (x.shape[0], 4 * 64 * 64 + 256) and cache.dtype == torch.float32
    assert torch.equal(cache[:, 4 * 64 * 64:], x[:, -1].to(torch.bfloat16 if m.do_bfloat16 else x.dtype).float())
    return whole, torch.cat(outs, 1), first, outs[0]


def test_streaming_equals_whole_utterance_cpu_oracle(monkeypatch):
    """Host logic on the CPU with the oracle as the scan (fp32 wrapper): pieces == whole to fp32 rounding."""
    w, m = _wrapper(False, "cpu")

    def wkv6(r, k, v, wd, u, s0, rev):
        y, sT = oracle.fwd(r, k, v, wd, u, s0, rev=rev, want_state=True, prec="f32")
        return y.to(r.dtype), sT.float()

    monkeypatch.setattr(w.autograd, "wkv6", wkv6)
    monkeypatch.setattr(w.autograd, "RUN_CUDA_RWKV6_FP32", lambda r, k, v, wd, u: wkv6(r, k, v, wd, u, None, False)[0])
    x = torch.randn(2, CUTS[-1], 256, generator=torch.Generator().manual_seed(12))
    whole, pieces, first, first_streamed = _stream_vs_whole(m, x)
    assert oracle.rel_err(pieces, whole) <= 1e-5
    assert oracle.rel_err(first_streamed, first) <= 1e-6      # a zero cache is the start of an utterance


@pytest.mark.gpu
@pytest.mark.parametrize("bf16", [False, True])
def test_streaming_equals_whole_utterance_gpu(bf16):
    """The same on the B200 kernels (fused prologue with the edge frame, scan with S_0 / S_T): fp32 1e-5, bf16 2e-2."""
    _, m = _wrapper(bf16, "cuda:0")
    x = torch.randn(2, CUTS[-1], 256, generator=torch.Generator().manual_seed(12)).cuda()
    whole, pieces, first, first_streamed = _stream_vs_whole(m, x)
    tol = 2e-2 if bf16 else 1e-5
    assert oracle.rel_err(pieces.cpu(), whole.cpu().double()) <= tol
    assert oracle.rel_err(first_streamed.cpu(), first.cpu().double()) <= tol
