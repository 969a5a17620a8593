"""Debug helper (GPU box): one RWKV_TmixWrapper_bidirectional block of this package against the reference's formulation on its
own CUDA kernels (oracle/_ref/libwkv6_ref.so), output by output."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
import paper_accurate_fast_cheap_b200.wrappers as W
ref = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libwkv6_ref.so"))
vp = lambda t: ctypes.c_void_p(t.data_ptr())
dev = "cuda:0"
torch.manual_seed(777)
m = W.RWKV_TmixWrapper_bidirectional(64, 512, 12, "v6", "bi", 4096, True, 3).to(dev).eval()
x = torch.randn(4, 200, 512, device=dev)
xb = x.to(torch.bfloat16)
f, b = m._blocks()
rel = lambda p, q: float((p.double() - q.double()).abs().max() / q.double().abs().max())
with torch.no_grad():
    for name, blk, rev in (("fwd", f, False), ("bwd", b, True)):
        r, k, v, w = blk.prologue(xb, rev)
        xin = torch.flip(xb, [1]) if rev else xb
        r2, k2, v2, w2 = blk.prologue(xin, False)
        if rev:
            r2, k2, v2, w2 = (torch.flip(t, [1]).contiguous() for t in (r2, k2, v2, w2))
        print(name, "prologue diffs", [rel(a, c) for a, c in zip((r, k, v, w), (r2, k2, v2, w2))])
        import paper_accurate_fast_cheap_b200 as pk
        y, _ = pk.wkv6_forward_raw(r, k, v, w, blk.time_faaaa, reverse=rev)
        rr, kk, vv, ww = ((torch.flip(t, [1]).contiguous() if rev else t) for t in (r, k, v, w))
        y2 = torch.empty_like(rr)
        rc = ref.wkv6_ref_forward(0, 4, 200, 512, 8, vp(rr), vp(kk), vp(vv), vp(ww), vp(blk.time_faaaa), vp(y2))
        torch.cuda.synchronize()
        if rev:
            y2 = torch.flip(y2, [1])
        print(name, "scan rc", rc, "diff", rel(y, y2), "max|y|", float(y.abs().max()), float(y2.abs().max()))
    out = m(x, x, x)[0]
    print("module out max", float(out.abs().max()))
