"""Debug helper (GPU box): clock64() phase trace of CTA 0, chunk 4, of the chunk kernels on the bench workload."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import paper_accurate_fast_cheap_b200 as pk
from paper_accurate_fast_cheap_b200 import _lib, ops
from test_gpu_parity import _inputs, _cuda

L = _lib.lib()
buf = (ctypes.c_ulonglong * 128)()
B, T, H = 16, 750, 8
a = _cuda(_inputs(B, T, H, torch.bfloat16, seed=1)); b = _cuda(_inputs(B, T, H, torch.bfloat16, seed=2, layer=7))
f, r_ = tuple(a[n] for n in "rkvwu"), tuple(b[n] for n in "rkvwu")
ck = ops.checkpoint_buffer(a["r"], 2)
for _ in range(3):
    pk.bi_forward_raw(f, r_, ckpt=ck); pk.bi_backward_raw(f, r_, a["gy"], b["gy"], ckpt=ck)
torch.cuda.synchronize()
L.wkv6_b200_debug_trace(1, None)
pk.bi_forward_raw(f, r_, ckpt=ck); pk.bi_backward_raw(f, r_, a["gy"], b["gy"], ckpt=ck)
torch.cuda.synchronize()
L.wkv6_b200_debug_trace(1, buf)
v = list(buf)
for name, base, n in (("fwd", 0, 15), ("bwd", 64, 16)):
    for who, off in (("t0 ", 0), ("t32", 16)):
        ts = v[base + off: base + off + n]
        t0 = ts[0]
        print(name, who, " ".join(f"{i}:{(x - t0) if x else -1}" for i, x in enumerate(ts)))
L.wkv6_b200_debug_trace(0, None)
