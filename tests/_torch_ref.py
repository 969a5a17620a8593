"""Plain PyTorch restatement of the WKV6 recurrence used to pin the
C oracle through torch.autograd.  Test helper only."""
import torch


def wkv6_recurrence(r, k, v, w, u, s0=None, rev=False):
    """r,k,v,w [B,T,C]; u [H,N]; s0 [B,H,N,N] (value i, key j).  Returns y [B,T,C], S_T."""
    B, T, C = r.shape
    H, N = u.shape
    rr, kk, vv, ww = (x.view(B, T, H, N) for x in (r, k, v, w))
    S = torch.zeros(B, H, N, N, dtype=r.dtype) if s0 is None else s0
    ys = [None] * T
    order = range(T - 1, -1, -1) if rev else range(T)
    for t in order:
        rt, kt, vt = rr[:, t], kk[:, t], vv[:, t]                 # [B,H,N]
        d = torch.exp(-torch.exp(ww[:, t]))                       # [B,H,N] over key j
        kv = vt.unsqueeze(-1) * kt.unsqueeze(-2)                  # [B,H,i,j]
        y = ((u.view(1, H, 1, N) * kv + S) * rt.unsqueeze(-2)).sum(-1)   # sum over j
        ys[t] = y.reshape(B, C)
        S = S * d.unsqueeze(-2) + kv
    return torch.stack(ys, 1), S


def make_inputs(B, T, H, N, seed=777, dtype=torch.float64, layer=0, n_layers=12, w_mode="init",
                with_state=False):
    """Synthetic op-level inputs of SURVEY.md section 8(d): r,k,v ~ N(0,1); w = time_decay init + 0.5 N(0,1)
    (model.py:248-251), the stress variant U(-8, 1.5) or the clamp variant (a third of the channels in U(1.7, 3.0)); u = time_faaaa init (model.py:256-260)."""
    C = H * N
    g = torch.Generator().manual_seed(seed)
    r, k, v, gy = (torch.randn(B, T, C, generator=g, dtype=torch.float64) for _ in range(4))
    ratio = layer / max(n_layers - 1, 1)
    n = torch.arange(C, dtype=torch.float64)
    if w_mode == "init":
        base = -6 + 5 * (n / max(C - 1, 1)) ** (0.7 + 1.3 * ratio)
        w = base.view(1, 1, C) + 0.5 * torch.randn(B, T, C, generator=g, dtype=torch.float64)
    elif w_mode == "stress":
        w = torch.rand(B, T, C, generator=g, dtype=torch.float64) * 9.5 - 8.0
    elif w_mode == "clamp":
        # the region where the chunk kernels' per-step decay floor is active (w > ln 5 = 1.609, csrc/wkv6_umma.cuh):
        # every third channel decays with w ~ U(1.7, 3.0), i.e. exp(-exp(w)) between 4e-3 and 2e-9; the rest: init
        base = -6 + 5 * (n / max(C - 1, 1)) ** (0.7 + 1.3 * ratio)
        w = base.view(1, 1, C) + 0.5 * torch.randn(B, T, C, generator=g, dtype=torch.float64)
        hot = 1.7 + 1.3 * torch.rand(B, T, C, generator=g, dtype=torch.float64)
        w = torch.where((torch.arange(C) % 3 == 1).view(1, 1, C), hot, w)
    else:
        raise ValueError(w_mode)
    zig = ((n + 1) % 3 - 1) * 0.1
    u = (ratio * (1 - n / max(C - 1, 1)) + zig).view(H, N)
    out = dict(r=r, k=k, v=v, w=w, u=u, gy=gy)
    if with_state:
        out["s0"] = torch.randn(B, H, N, N, generator=g, dtype=torch.float64)
        out["gsT"] = torch.randn(B, H, N, N, generator=g, dtype=torch.float64)
    return {kk_: vv_.to(dtype) for kk_, vv_ in out.items()}


def tmix_shift(x, reverse=False, x_edge=None):
    """xx = shift(x) - x of RWKV_Tmix_x060c.forward (wenet/rwkv_v6/src/model.py:272-273: nn.ZeroPad2d((0,0,1,-1)));
    `reverse`: the flip-free right-to-left block looks at t+1 instead; `x_edge` [B,C]: the frame
    beyond the edge instead of the zero padding (streaming)."""
    import torch.nn.functional as F
    if x_edge is not None:
        e = x_edge.view(x.shape[0], 1, x.shape[2])
        return (torch.cat([x[:, 1:], e], 1) if reverse else torch.cat([e, x[:, :-1]], 1)) - x
    if reverse:
        return F.pad(x[:, 1:], (0, 0, 0, 1)) - x
    return F.pad(x[:, :-1], (0, 0, 1, 0)) - x


def tmix_mix_stage1(x, maa_x, reverse=False, x_edge=None):
    """model.py:275: xxx = x + xx * time_maa_x (eager ops in x's dtype)."""
    return x + tmix_shift(x, reverse, x_edge) * maa_x


def tmix_mix_stage2(x, m, maa_r, maa_k, maa_v, maa_w, reverse=False, x_edge=None):
    """model.py:280-287: x + xx * (time_maa_c + m_c) for c in r, k, v, w; m is [4,B,T,C]."""
    xx = tmix_shift(x, reverse, x_edge)
    return tuple(x + xx * (maa + mc) for maa, mc in zip((maa_r, maa_k, maa_v, maa_w), m.unbind(0)))
