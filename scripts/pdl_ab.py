"""Back-to-back forward / backward launches on a side stream and on the default stream (A/B helper for the
programmatic-dependent-launch attribute: run once as is and once with WKV6_NO_PDL=1)."""
import os
import sys
import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paper_accurate_fast_cheap_b200 as pk
from paper_accurate_fast_cheap_b200 import ops

B, T, H, N = 16, 750, 8, 64
C = H * N
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(1)


def mk(nset):
    out = []
    for _ in range(nset):
        d = []
        for _dir in range(2):
            r, k, v = (torch.randn(B, T, C, device=dev, generator=g).bfloat16() for _ in range(3))
            w = (-6 + 5 * torch.rand(B, T, C, device=dev, generator=g)).bfloat16()
            u = torch.randn(C, device=dev, generator=g).bfloat16() * 0.3
            d.append((r, k, v, w, u))
        out.append(d)
    return out


sets = mk(12)
gy = torch.randn(B, T, C, device=dev, generator=g).bfloat16()
ck = ops.checkpoint_buffer(sets[0][0][0], 2)


def run(stream, what, reps=20):
    with torch.cuda.stream(stream):
        def once():
            for f, b in sets:
                if what == "fwd":
                    pk.bi_forward_raw(f, b)
                else:
                    pk.bi_backward_raw(f, b, gy, gy, ckpt=ck)
        if what == "bwd":
            pk.bi_forward_raw(sets[0][0], sets[0][1], ckpt=ck)
        for _ in range(3):
            once()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        stream.synchronize()
        e0.record(stream)
        for _ in range(reps):
            once()
        e1.record(stream)
        stream.synchronize()
        return e0.elapsed_time(e1) * 1e3 / (reps * len(sets))


side = torch.cuda.Stream()
tag = "NO_PDL" if os.environ.get("WKV6_NO_PDL") else "PDL"
for what in ("fwd", "bwd"):
    print(f"{tag} {what}: default stream {run(torch.cuda.current_stream(), what):.2f} us/launch, side stream {run(side, what):.2f} us/launch")
