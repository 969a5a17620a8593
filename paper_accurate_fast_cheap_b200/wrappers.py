"""Host-side mirror of the reference's attention-slot modules for the WKV6 path.

Same class names, constructor signatures, parameter / state-dict names, forward signatures and return values as

    wenet/rwkv_v6/src/model.py:217-325                                   RWKV_Tmix_x060c
    wenet/rwkv_v6/rwkv_wrapper.py:5-83                                   RWKV_TmixWrapper
    wenet/rwkv_v6/rwkv_wrapper_bidirectional.py:7-64                     RWKV_TmixWrapper_bidirectional
    wenet/rwkv_v6/rwkv_wrapper_bidirectional_direction_dropout.py:7-97   ..._direction_dropout
    wenet/rwkv_v6/rwkv_wrapper_bidirectional_direction_dropout_both.py   ..._direction_dropout_both

so a checkpoint of the reference loads unchanged and `install()` can swap them into wenet's
`WENET_ATTENTION_CLASSES` (wenet/utils/class_utils.py:82-86).  What is different is HOW the block runs:

* the right-to-left direction never flips anything: its token shift is `x[t+1] - x[t]`, every
  other prologue/epilogue op is position-wise, and the scan itself walks t = T-1..0 inside the kernel;
* both directions' scans are ONE launch (`bi_wkv6`), direction dropout becomes the launch's `flags`;
* `(fwd + bwd) / 2` (+ the cast to float) is one fused pass (`combine`).

The GEMMs / LayerNorm around the scan stay torch ops ("surrounds the op, runs unchanged").
No CPU path: the scan raises on CPU tensors.
"""
from __future__ import annotations

import os
from typing import Optional, Tuple

import torch
import torch.nn as nn
import torch.nn.functional as F

from . import _lib, autograd, ops


class RWKV_Tmix_x060c(nn.Module):
    """Parameters, initialisation order (RNG-identical) and math of model.py:217-325."""

    def __init__(self, head_size: int, n_layers: int, n_embd: int, dim_att: int, layer_id: int):
        super().__init__()
        self.layer_id, self.head_size, self.n_head = layer_id, head_size, dim_att // head_size
        self.dim_att, self.n_embd, self.n_layers = dim_att, n_embd, n_layers
        assert dim_att % self.n_head == 0
        with torch.no_grad():
            ratio_0_to_1 = layer_id / (n_layers - 1)
            ratio_1_to_almost0 = 1.0 - (layer_id / n_layers)
            ddd = (torch.arange(n_embd, dtype=torch.float64) / n_embd).float().view(1, 1, n_embd)
            self.time_maa_x = nn.Parameter(1.0 - torch.pow(ddd, ratio_1_to_almost0))
            self.time_maa_r = nn.Parameter(1.0 - torch.pow(ddd, 0.5 * ratio_1_to_almost0))
            self.time_maa_k = nn.Parameter(1.0 - torch.pow(ddd, ratio_1_to_almost0))
            self.time_maa_v = nn.Parameter(1.0 - (torch.pow(ddd, ratio_1_to_almost0) + 0.3 * ratio_0_to_1))
            self.time_maa_w = nn.Parameter(1.0 - torch.pow(ddd, ratio_1_to_almost0))
            self.time_maa_rkvw_w1 = nn.Parameter(torch.zeros(n_embd, 32 * 4))
            self.time_maa_rkvw_w2 = nn.Parameter(torch.zeros(4, 32, n_embd).uniform_(-0.01, 0.01))
            n = torch.arange(dim_att, dtype=torch.float64)
            decay_speed = (-6 + 5 * (n / (dim_att - 1)) ** (0.7 + 1.3 * ratio_0_to_1)).float()
            self.time_decay = nn.Parameter(decay_speed.reshape(1, 1, dim_att))
            self.time_decay_w1 = nn.Parameter(torch.zeros(n_embd, 64))
            self.time_decay_w2 = nn.Parameter(torch.zeros(64, dim_att).uniform_(-0.01, 0.01))
            zigzag = ((n + 1) % 3 - 1) * 0.1
            faaaa = (ratio_0_to_1 * (1 - (n / (dim_att - 1))) + zigzag).float()
            self.time_faaaa = nn.Parameter(faaaa.reshape(self.n_head, head_size))
        self.receptance = nn.Linear(n_embd, dim_att, bias=False)
        self.key = nn.Linear(n_embd, dim_att, bias=False)
        self.value = nn.Linear(n_embd, dim_att, bias=False)
        self.output = nn.Linear(dim_att, n_embd, bias=False)
        self.ln_x = nn.LayerNorm(dim_att)

    def prologue(self, x: torch.Tensor, reverse: bool = False, x_edge: Optional[torch.Tensor] = None):
        """x [B,T,C] -> contiguous r, k, v, w (model.py:270-289).  `reverse`: the token shift looks at t+1.
        `x_edge` [B,C]: the frame beyond the edge the shift looks across (streaming: last frame of the previous piece);
        None = zeros, as at the start of an utterance."""
        B, T, C = x.size()
        if x.is_cuda and x.dtype in (torch.bfloat16, torch.float32) and C % 8 == 0:
            # the elementwise stages run fused (forward bit-exact with the eager ops below; with autograd on, their backward
            # is one kernel per stage as well: ops.TmixStage1 / TmixStage2), the LoRA GEMMs in between are torch ops
            xr, xk, xv, ww = ops.tmix_mix_autograd(x, self.time_maa_x,
                                                   (self.time_maa_r, self.time_maa_k, self.time_maa_v, self.time_maa_w),
                                                   self.time_maa_rkvw_w1, self.time_maa_rkvw_w2, reverse, x_edge)
            w = self.time_decay + torch.tanh(ww @ self.time_decay_w1) @ self.time_decay_w2
            return (self.receptance(xr).contiguous(), self.key(xk).contiguous(), self.value(xv).contiguous(),
                    w.contiguous())
        if x_edge is not None:
            e = x_edge.to(x.dtype).view(B, 1, C)
            xx = (torch.cat([x[:, 1:], e], 1) if reverse else torch.cat([e, x[:, :-1]], 1)) - x
        elif reverse:
            xx = F.pad(x[:, 1:], (0, 0, 0, 1)) - x
        else:
            xx = F.pad(x[:, :-1], (0, 0, 1, 0)) - x                      # nn.ZeroPad2d((0,0,1,-1))
        xxx = x + xx * self.time_maa_x
        xxx = torch.tanh(xxx @ self.time_maa_rkvw_w1).view(B * T, 4, -1).transpose(0, 1)
        xxx = torch.bmm(xxx, self.time_maa_rkvw_w2).view(4, B, T, C)
        mr, mk, mv, mw = xxx.unbind(dim=0)
        r = self.receptance(x + xx * (self.time_maa_r + mr))
        k = self.key(x + xx * (self.time_maa_k + mk))
        v = self.value(x + xx * (self.time_maa_v + mv))
        ww = x + xx * (self.time_maa_w + mw)
        w = self.time_decay + torch.tanh(ww @ self.time_decay_w1) @ self.time_decay_w2
        return r.contiguous(), k.contiguous(), v.contiguous(), w.contiguous()

    def epilogue(self, y: torch.Tensor) -> torch.Tensor:
        return self.output(self.ln_x(y))                                  # model.py:323-324

    def forward(self, x: torch.Tensor, reverse: bool = False) -> torch.Tensor:
        r, k, v, w = self.prologue(x, reverse)
        if reverse:
            y, _ = autograd.wkv6(r, k, v, w, self.time_faaaa, None, True)
        elif x.dtype == torch.float32:
            y = autograd.RUN_CUDA_RWKV6_FP32(r, k, v, w, self.time_faaaa)     # model.py:296-299
        else:
            y = autograd.RUN_CUDA_RWKV6(r, k, v, w, self.time_faaaa)
        return self.epilogue(y)

    def forward_stream(self, x: torch.Tensor, state: Optional[torch.Tensor], x_last: Optional[torch.Tensor]):
        """One piece of a left-to-right stream: `state` [B,H,N,N] fp32 is the WKV state after the previous
        piece, `x_last` [B,C] its last input frame (None, None at the start of an utterance).  Returns (out, new_state,
        new_x_last); the pieces' outputs concatenated equal the output of one call over the whole utterance."""
        r, k, v, w = self.prologue(x, False, x_last)
        y, s_new = autograd.wkv6(r, k, v, w, self.time_faaaa, state, False)
        return self.epilogue(y), s_new, x[:, -1]


_EMPTY_MASK = torch.ones((0, 0, 0), dtype=torch.bool)
_EMPTY_POS = torch.empty(0)
_EMPTY_CACHE = torch.zeros((0, 0, 0, 0))


class RWKV_TmixWrapper(nn.Module):
    """Unidirectional attention-slot wrapper (rwkv_wrapper.py:5-83): MHA-shaped signature, cache passed through."""

    def __init__(self, head_size: int, dim_att: int, num_blocks: int, rnn_att_version: Optional[str] = None,
                 rnn_att_direction: Optional[str] = None, ctx_len: int = 2048, do_bfloat16: bool = True,
                 layer_id: int = 1):
        super().__init__()
        self.head_size, self.dim_att, self.num_blocks = head_size, dim_att, num_blocks
        self.rnn_att_version, self.rnn_att_direction = rnn_att_version, rnn_att_direction
        self.ctx_len, self.do_bfloat16, self.layer_id = ctx_len, do_bfloat16, layer_id      # ctx_len: no T limit here
        self.n_head, self.n_embd = dim_att // head_size, dim_att
        assert head_size == ops.HEAD_SIZE, "this build supports head_size 64 (every shipped config)"
        self.tmix_block = RWKV_Tmix_x060c(head_size=head_size, n_layers=num_blocks, n_embd=dim_att, dim_att=dim_att,
                                          layer_id=layer_id)
        if do_bfloat16:
            self.tmix_block = self.tmix_block.to(dtype=torch.bfloat16)

    def forward(self, query, key=None, value=None, mask=_EMPTY_MASK, pos_emb=_EMPTY_POS, cache=_EMPTY_CACHE
                ) -> Tuple[torch.Tensor, torch.Tensor]:
        qd = query.dtype
        x = query.to(dtype=torch.bfloat16) if self.do_bfloat16 else query
        if cache.numel() > 0:
            # streaming (the reference passes the cache through untouched, rwkv_wrapper.py:81): the cache
            # carries the WKV state and the last input frame, see init_cache()
            B, C, nn_ = x.shape[0], self.n_embd, self.n_head * self.head_size * self.head_size
            assert cache.shape == (B, nn_ + C) and cache.dtype == torch.float32, "cache: see RWKV_TmixWrapper.init_cache"
            state = cache[:, :nn_].reshape(B, self.n_head, self.head_size, self.head_size).contiguous()
            out, s_new, x_last = self.tmix_block.forward_stream(x, state, cache[:, nn_:].to(x.dtype).contiguous())
            new_cache = torch.cat([s_new.reshape(B, nn_).float(), x_last.float()], 1)
            return (out.to(dtype=qd) if self.do_bfloat16 else out), new_cache
        out = self.tmix_block(x)
        if self.do_bfloat16:
            out = out.to(dtype=qd)
        return out, cache

    def init_cache(self, batch: int, device=None) -> torch.Tensor:
        """Cache of a stream that has seen nothing yet: [B, H*N*N + C] fp32 = WKV state [B,H,N,N] (zeros) followed by the
        last input frame (zeros: the token shift of frame 0 sees zero, model.py:272).  Pass it as `cache`; every call
        returns the cache for the next piece."""
        return torch.zeros(batch, self.n_head * self.head_size * self.head_size + self.n_embd, dtype=torch.float32, device=device)


def _bi_scan(fwd: RWKV_Tmix_x060c, bwd: RWKV_Tmix_x060c, x: torch.Tensor, run_l2r: bool, run_r2l: bool):
    """Both Tmix blocks on the same un-flipped x: prologues and ONE scan launch.  Returns (y_f, y_b), None where dropped."""
    set_f = (*fwd.prologue(x, reverse=False), fwd.time_faaaa) if run_l2r else None
    set_b = (*bwd.prologue(x, reverse=True), bwd.time_faaaa) if run_r2l else None
    return autograd.bi_wkv6(set_f if run_l2r else set_b, set_b if run_r2l else set_f, run_l2r, run_r2l)


def _bi_block(fwd: RWKV_Tmix_x060c, bwd: RWKV_Tmix_x060c, x: torch.Tensor, run_l2r: bool, run_r2l: bool):
    """_bi_scan + the two epilogues.  Returns (out_f, out_b) in x's dtype at natural time positions; None for a direction
    that did not run."""
    y_f, y_b = _bi_scan(fwd, bwd, x, run_l2r, run_r2l)
    return (fwd.epilogue(y_f) if run_l2r else None), (bwd.epilogue(y_b) if run_r2l else None)


def _fused_epilogue_ok(x: torch.Tensor) -> bool:
    return x.is_cuda and not torch.is_grad_enabled() and x.dtype in (torch.bfloat16, torch.float32) and x.shape[-1] % 8 == 0


def _bi_block_mean(owner: nn.Module, fwd: RWKV_Tmix_x060c, bwd: RWKV_Tmix_x060c, x: torch.Tensor) -> torch.Tensor:
    """(F(x) + G(x)) / 2 of both directions with the epilogue as ONE GEMM (inference): the two ln_x write side by side
    (ops.ln_pair_cat) and [ln_f(y_f) || ln_b(y_b)] x [W_f ; W_b]^T / 2 replaces two output projections, the add and the
    halving (model.py:323-324 twice + rwkv_wrapper_bidirectional.py:48).  Result in x's dtype; the sum is formed in the
    GEMM's fp32 accumulator and rounded once (the reference rounds each projection, then the sum)."""
    y_f, y_b = _bi_scan(fwd, bwd, x, True, True)
    z = ops.ln_pair_cat(y_f, y_b, fwd.ln_x, bwd.ln_x)
    wf, wb = fwd.output.weight, bwd.output.weight
    key = (wf.data_ptr(), wf._version, wb.data_ptr(), wb._version, wf.dtype, wf.device)
    cached = getattr(owner, "_wcat_half", None)
    if cached is None or cached[0] != key:
        cached = (key, (torch.cat([wf.detach(), wb.detach()], dim=1) * 0.5).contiguous())     # [C_out, 2C]; x 0.5 is exact
        owner._wcat_half = cached
    return F.linear(z, cached[1])


class _BiBase(nn.Module):
    def _make(self, head_size, dim_att, num_blocks, ver, direction, ctx_len, do_bfloat16, layer_id):
        self.rwkv_wrapper_forward = RWKV_TmixWrapper(head_size, dim_att, num_blocks, ver, direction, ctx_len,
                                                     do_bfloat16, layer_id)
        self.rwkv_wrapper_backward = RWKV_TmixWrapper(head_size, dim_att, num_blocks, ver, direction, ctx_len,
                                                      do_bfloat16, layer_id)

    def _blocks(self):
        return self.rwkv_wrapper_forward.tmix_block, self.rwkv_wrapper_backward.tmix_block


class RWKV_TmixWrapper_bidirectional(_BiBase):
    """out = (F(x) + flip(G(flip(x)))) / 2, .float()  (rwkv_wrapper_bidirectional.py:30-64), flip-free."""

    def __init__(self, head_size: int, dim_att: int, num_blocks: int, rnn_att_version: str = None,
                 rnn_att_direction: str = None, ctx_len: int = 2048, do_bfloat16: bool = True, layer_id: int = 1):
        super().__init__()
        self.do_bfloat16 = do_bfloat16
        self._make(head_size, dim_att, num_blocks, rnn_att_version, rnn_att_direction, ctx_len, do_bfloat16, layer_id)
        self.rwkv_wrapper_forward.do_bfloat16 = False          # conversion handled here, as in the reference (:25-28)
        self.rwkv_wrapper_backward.do_bfloat16 = False

    def forward(self, query, key=None, value=None, mask=_EMPTY_MASK, pos_emb=_EMPTY_POS, cache=_EMPTY_CACHE):
        x = query.to(dtype=torch.bfloat16) if self.do_bfloat16 else query
        if _fused_epilogue_ok(x):
            out = _bi_block_mean(self, *self._blocks(), x)
            return (out.float() if self.do_bfloat16 else out.to(query.dtype)), cache
        o_f, o_b = _bi_block(*self._blocks(), x, True, True)
        # the reference adds and halves in x's dtype, then casts (:48-54): round_sum reproduces the bf16 rounding
        out = autograd.combine(o_f, o_b, 0.5, round_sum=self.do_bfloat16, out_float=True)
        if not self.do_bfloat16:
            out = out.to(query.dtype)
        return out, cache

    forward_parallel = forward                                  # :66-116: same result, two streams in the reference


class RWKV_TmixWrapper_bidirectional_direction_dropout(_BiBase):
    """Training: L2R always, R2L kept with p = 0.8 (one CPU Bernoulli draw per call, ..._dropout.py:59-69).
    Eval: bidirectional / alternating / L2R-only by RWKV_BIDIRECTIONAL_LAYERS and RWKV_ALT_DECODING (:22-33,70-91)."""
    BOTH = False

    def __init__(self, head_size: int, dim_att: int, num_blocks: int, rnn_att_version: str = None,
                 rnn_att_direction: str = None, ctx_len: int = 2048, do_bfloat16: bool = True, layer_id: int = 1):
        super().__init__()
        self.layer_id = layer_id
        self.do_bfloat16 = do_bfloat16
        self.bi_layers_actives = []
        self.bi_active = True
        self.alt_decoding = os.environ.get("RWKV_ALT_DECODING", "0") == "1"
        if os.environ.get("RWKV_BIDIRECTIONAL_LAYERS"):
            self.bi_layers_actives = [int(v) for v in os.environ.get("RWKV_BIDIRECTIONAL_LAYERS").split(",")]
            if layer_id not in self.bi_layers_actives:
                self.bi_active = False
        self._make(head_size, dim_att, num_blocks, rnn_att_version, rnn_att_direction, ctx_len, do_bfloat16, layer_id)

    def _directions(self) -> Tuple[bool, bool]:
        if self.training:
            p = 0.2
            keep = torch.distributions.bernoulli.Bernoulli(1 - p).sample((1,))
            if keep == 1:
                return True, True
            if not self.BOTH:
                return True, False
            choose = torch.distributions.bernoulli.Bernoulli(1 - p).sample((1,))     # ..._both.py:63-71
            return (True, False) if choose <= 0.5 else (False, True)
        if self.bi_active:
            return True, True
        if self.alt_decoding:
            return (True, False) if self.layer_id % 2 == 0 else (False, True)
        return True, False

    def forward(self, query, key=None, value=None, mask=_EMPTY_MASK, pos_emb=_EMPTY_POS, cache=_EMPTY_CACHE):
        run_f, run_b = self._directions()
        return self.forward_directions(query, run_f, run_b, cache)

    def forward_directions(self, query, run_f: bool, run_b: bool, cache=_EMPTY_CACHE):
        """forward() with the direction outcome given instead of drawn (GraphedTraining captures one graph per outcome)."""
        qd = query.dtype
        x = query.to(dtype=torch.bfloat16) if self.do_bfloat16 else query      # inner wrappers cast per direction
        if run_f and run_b and _fused_epilogue_ok(x):
            return _bi_block_mean(self, *self._blocks(), x).to(qd), cache
        o_f, o_b = _bi_block(*self._blocks(), x, run_f, run_b)
        if run_f and run_b:
            # inner wrappers hand back query-dtype tensors; the reference averages those (fp32 when the encoder is)
            out = autograd.combine(o_f, o_b, 0.5, round_sum=False, out_float=(qd == torch.float32)).to(qd)
        else:
            out = (o_f if run_f else o_b).to(qd)
        return out, cache


class RWKV_TmixWrapper_bidirectional_direction_dropout_both(RWKV_TmixWrapper_bidirectional_direction_dropout):
    """As above, but a dropped step keeps L2R w.p. 0.2 and R2L w.p. 0.8 (the `choose <= 0.5` quirk)."""
    BOTH = True


# the five keys of WENET_ATTENTION_CLASSES this path owns (wenet/utils/class_utils.py:82-86)
WENET_ATTENTION_CLASSES_B200 = {
    "rwkv_tmix60": RWKV_TmixWrapper,
    "rwkv_tmix60_bidirectional": RWKV_TmixWrapper_bidirectional,
    "rwkv_tmix60_bidirectional2": RWKV_TmixWrapper_bidirectional,
    "rwkv_tmix60_dir_layer_drop": RWKV_TmixWrapper_bidirectional_direction_dropout,
    "rwkv_tmix60_dir_layer_drop_both": RWKV_TmixWrapper_bidirectional_direction_dropout_both,
}


def install_ops() -> None:
    """Makes the UNMODIFIED reference model.py use these kernels: registers torch.ops.wkv6.* (same schemas as
    wkv6_op.cpp:9-41) and turns the reference's JIT build of its own `wkv6` extension (model.py:105-106) into a
    no-op -- a second TORCH_LIBRARY(wkv6) would abort.  Call before importing wenet.rwkv_v6.src.model."""
    _lib.load_torch_ops()
    from torch.utils import cpp_extension
    if getattr(cpp_extension.load, "_wkv6_b200", False):
        return
    orig = cpp_extension.load

    def load(name, *a, **kw):
        if name == "wkv6":
            return None
        return orig(name, *a, **kw)

    load._wkv6_b200 = True
    cpp_extension.load = load


def install(class_utils=None) -> None:
    """install_ops() + swap the five attention classes in wenet.utils.class_utils (if given or importable)."""
    install_ops()
    if class_utils is None:
        try:
            from wenet.utils import class_utils as class_utils     # noqa: PLC0415
        except Exception:
            return
    class_utils.WENET_ATTENTION_CLASSES.update(WENET_ATTENTION_CLASSES_B200)


class GraphedInference:
    """Replays the attention-slot modules of an encoder as ONE CUDA graph (inference, fixed [B, T, C] shape).

    Eagerly, a bidirectional block is ~45 launches of which the scan is one: at configs[1]'s size the host cannot issue
    them as fast as the B200 retires them (7.1 ms per 12 blocks against ~2 ms of kernel time).  Every entry point of
    libwkv6_b200.so is capturable (no synchronisation, no allocation; launches go to the current stream), so the whole
    stack is captured once and replayed:

        g = GraphedInference(modules, example_x)      # modules: list of RWKV_TmixWrapper* in eval mode
        outs = g(x)                                   # list of per-module outputs for this x (static buffers: copy to keep)

    `chain=True` feeds every module the previous module's output (as a bare stack of blocks); the default gives every
    module the same x, which is how an encoder layer calls it (its feed-forward / convolution modules run in between).
    """

    def __init__(self, modules, example_x: torch.Tensor, chain: bool = False, warmup: int = 2):
        assert example_x.is_cuda, "GraphedInference: CUDA tensors only"
        self.modules, self.chain = list(modules), chain
        self.x = example_x.clone()
        side = torch.cuda.Stream(device=example_x.device)
        side.wait_stream(torch.cuda.current_stream(example_x.device))
        with torch.cuda.stream(side), torch.no_grad():
            for _ in range(warmup):                    # first calls opt in to shared memory sizes, build cuBLAS handles, ...
                self._run()
        torch.cuda.current_stream(example_x.device).wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.no_grad(), torch.cuda.graph(self.graph):
            self.outs = self._run()

    def _run(self):
        outs, h = [], self.x
        for m in self.modules:
            o = m(h, h, h)[0]
            outs.append(o)
            if self.chain:
                h = o
        return outs

    def __call__(self, x: torch.Tensor):
        self.x.copy_(x)
        self.graph.replay()
        return self.outs


class _FixedDirections(nn.Module):
    """x -> out of an attention-slot module with a fixed direction outcome (no host-side draw inside: capturable)."""

    def __init__(self, module: nn.Module, run_f: bool, run_b: bool):
        super().__init__()
        self.module, self.run_f, self.run_b = module, run_f, run_b

    def forward(self, x):
        if hasattr(self.module, "forward_directions"):
            return self.module.forward_directions(x, self.run_f, self.run_b)[0]
        return self.module(x, x, x)[0]


class GraphedTraining:
    """Training step of ONE attention-slot module with forward and backward as CUDA graphs (fixed [B, T, C] shape).

    The autograd Functions of this package (scan, fused prologue stages, combine) only launch kernels of libwkv6_b200.so and
    torch ops on the current stream, so `torch.cuda.make_graphed_callables` captures a module's forward and its backward;
    the eager training step is host-bound at configs[3]'s size (27 ms per 12 blocks against ~4 ms of kernel time, 11 ms
    graphed).  Direction dropout is a HOST decision (one CPU Bernoulli draw per call, in the reference's order): one pair
    of graphs per outcome -- both directions, L2R only, R2L only -- is captured when the object is built, and every call
    draws exactly like the eager module and replays that outcome's graphs; a dropped direction's parameters get
    no gradient (`allow_unused_input`), as in eager mode.  In eval mode the module is called directly.

        slot = GraphedTraining(module, example_x)        # module.train(); example_x: [B, T, C] on the GPU
        out = slot(x); loss(out).backward()              # out is a static buffer of the outcome's graph
    """

    def __init__(self, module: nn.Module, example_x: torch.Tensor, warmup: int = 3):
        assert example_x.is_cuda, "GraphedTraining: CUDA tensors only"
        self.module, self.warmup = module, warmup
        self._example = example_x.detach().clone()
        self._graphs = {}
        # all outcomes are captured up front: a capture AFTER a replayed backward of another outcome's graph fails (the
        # autograd engine then makes the legacy stream wait on the capturing stream for AccumulateGrad nodes that a live
        # graph keeps alive)
        if module.training:
            outcomes = ((True, True), (True, False), (False, True)) if hasattr(module, "_directions") else ((True, True),)
            for run in outcomes:
                self._capture(run)

    def _capture(self, run):
        fixed = _FixedDirections(self.module, *run)
        x = self._example.clone().requires_grad_(True)
        g = torch.cuda.make_graphed_callables(fixed, (x,), num_warmup_iters=self.warmup, allow_unused_input=True)
        self.module.zero_grad(set_to_none=True)          # the warm-up iterations ran real backward passes
        self._graphs[run] = g
        return g

    def __call__(self, x: torch.Tensor) -> torch.Tensor:
        m = self.module
        if not m.training:
            return m(x, x, x)[0]
        run = tuple(m._directions()) if hasattr(m, "_directions") else (True, True)
        g = self._graphs.get(run) or self._capture(run)
        return g(x)
