"""Times the fused token-shift / maa-mix stages (csrc/wkv6_tmix.cu) on the bench shape (B=16, T=750, C=512, bf16, 12
distinct blocks so that every launch reads cold data: 110 MB per stage-2 launch) against the eager torch ops they
replace, and prints one JSON line with the algorithmic bandwidth against the measured HBM peak."""
import ctypes
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests"))
import _torch_ref as ref
from paper_accurate_fast_cheap_b200 import _lib, ops

B, T, C, NL = 16, 750, 512, 12
dev = torch.device("cuda:0")
dt = torch.bfloat16
g = torch.Generator(device=dev).manual_seed(3)
xs = [torch.randn(B, T, C, device=dev, generator=g).to(dt) for _ in range(NL)]
ms = [(0.3 * torch.randn(4, B, T, C, device=dev, generator=g)).to(dt) for _ in range(NL)]
maa = [torch.rand(C, device=dev, generator=g).to(dt) for _ in range(5)]
maa3 = [t.view(1, 1, C) for t in maa]
out1 = [torch.empty_like(xs[0]) for _ in range(NL)]
out2 = [[torch.empty_like(xs[0]) for _ in range(4)] for _ in range(NL)]
L = _lib.lib()
p = lambda t: ctypes.c_void_p(t.data_ptr())



def st_now():
    return ctypes.c_void_p(torch.cuda.current_stream().cuda_stream)


def fused1():
    for l in range(NL):
        _lib.check(L.wkv6_b200_tmix_mix(0, B, T, C, p(xs[l]), None, None, p(maa[0]), None, None, None, None, 0, p(out1[l]), None, None, None, None, st_now()))


def fused2():
    for l in range(NL):
        _lib.check(L.wkv6_b200_tmix_mix(0, B, T, C, p(xs[l]), None, p(ms[l]), None, p(maa[1]), p(maa[2]), p(maa[3]), p(maa[4]), 0, None,
                                        *[p(o) for o in out2[l]], st_now()))


def eager1():
    for l in range(NL):
        ref.tmix_mix_stage1(xs[l], maa3[0])


def eager2():
    for l in range(NL):
        ref.tmix_mix_stage2(xs[l], ms[l], maa3[1], maa3[2], maa3[3], maa3[4])


def timed(fn, reps=20):
    """us per block.  The 12 launches of a pass are captured in a CUDA graph and the graph is replayed: a stage-1 launch
    is ~5 us of GPU work, less than the ~10 us a ctypes call costs on the host, so timing eager launches would time Python."""
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        for _ in range(3):
            fn()
    torch.cuda.current_stream().wait_stream(side)
    torch.cuda.synchronize()
    g = torch.cuda.CUDAGraph()
    with torch.cuda.graph(g):
        fn()
    g.replay()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        g.replay()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / (reps * NL) * 1e3


peak = 6449.4
try:
    peak = json.load(open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "MEASURED_PEAKS.json")))["hbm_gbs"]
except Exception:
    pass
n = B * T * C * 2
res = {}
# row a7: (y_f + y_b) / 2 -> fp32 (rwkv_wrapper_bidirectional.py:48-56): 2 bf16 reads + 1 fp32 write = 8 bytes per element
ya = [torch.randn(B, T, C, device=dev, generator=g).to(dt) for _ in range(NL)]
yb = [torch.randn(B, T, C, device=dev, generator=g).to(dt) for _ in range(NL)]
yo = [torch.empty(B, T, C, device=dev, dtype=torch.float32) for _ in range(NL)]


def combine_fused():
    for l in range(NL):
        _lib.check(L.wkv6_b200_combine(0, ya[l].numel(), p(ya[l]), p(yb[l]), ctypes.c_float(0.5), 1, p(yo[l]), st_now()))


def combine_eager():
    for l in range(NL):
        ((ya[l] + yb[l]) / 2).float()


only = os.environ.get("TMIX_ONLY")          # e.g. stage2_fused: the one variant an ncu capture wants
for name, fn, passes in (("stage1_fused", fused1, 2), ("stage1_eager", eager1, 2), ("stage2_fused", fused2, 9), ("stage2_eager", eager2, 9),
                         ("combine_fused", combine_fused, 4), ("combine_eager", combine_eager, 4)):
    if only and name != only:
        continue
    us = timed(fn)
    res[name] = {"us_per_block": us, "algorithmic_GBps": passes * n / us / 1e3, "frac_of_hbm_peak": passes * n / us / 1e3 / peak}
print(json.dumps({"workload": "tmix token-shift + maa mixes, B=16 x T=750 x C=512 bf16, 12 distinct blocks", "hbm_peak_GBps": peak, **res}))
