"""A/B (GPU box): does mixing the two direction instantiations on one SM cost anything?  256 CTAs either as
B=16 x 8 heads x 2 directions (one bidirectional launch) or B=32 x 8 heads, one direction (forward-only / reverse-only)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import paper_accurate_fast_cheap_b200 as pk
from test_gpu_parity import _inputs, _cuda

T, H = 750, 8
sets16 = [_cuda(_inputs(16, T, H, torch.bfloat16, seed=s)) for s in range(8)]
sets32 = [_cuda(_inputs(32, T, H, torch.bfloat16, seed=20 + s)) for s in range(4)]

def timeit(fn, n=24):
    for i in range(4):
        fn(i)
    torch.cuda.synchronize()
    ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(n)]
    for i, (a, b) in enumerate(ev):
        a.record(); fn(i); b.record()
    torch.cuda.synchronize()
    ts = sorted(a.elapsed_time(b) for a, b in ev)
    return 1e3 * ts[len(ts) // 2]

def bi(i):
    a, b = sets16[(2 * i) % 8], sets16[(2 * i + 1) % 8]
    pk.bi_forward_raw(tuple(a[n] for n in "rkvwu"), tuple(b[n] for n in "rkvwu"))

def uni(rev):
    def f(i):
        a = sets32[i % 4]
        pk.wkv6_forward_raw(a["r"], a["k"], a["v"], a["w"], a["u"], reverse=rev)
    return f

print("bidirectional B=16 (mixed instantiations): %.1f us" % timeit(bi))
print("forward only  B=32                       : %.1f us" % timeit(uni(False)))
print("reverse only  B=32                       : %.1f us" % timeit(uni(True)))
