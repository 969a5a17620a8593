"""Long-form shapes of BASELINE.json configs[2] (one GPU's share at 8 / 4 / 2 / 1 GPUs: B = 1, 2, 4, 8 utterances of
T = 30 000 subsampled frames, C = 512, 8 heads, bf16, both directions, 12 blocks): frames/s of the bidirectional forward
with and without the time-segmented launch (csrc/wkv6_chunk_lean.cu).  Prints one JSON line."""
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import paper_accurate_fast_cheap_b200 as pk
from paper_accurate_fast_cheap_b200 import ops

T, H, N, NL = 30000, 8, 64, 12
C = H * N
dev = torch.device("cuda:0")
g = torch.Generator(device=dev).manual_seed(2)


def mk(B):
    r, k, v = (torch.randn(B, T, C, device=dev, generator=g).bfloat16() for _ in range(3))
    w = (-6 + 5 * torch.rand(B, T, C, device=dev, generator=g)).bfloat16()
    u = (0.3 * torch.randn(H, N, device=dev, generator=g)).bfloat16()
    return r, k, v, w, u


def step_ms(f, b, reps=3):
    def once():
        for _ in range(NL):
            pk.bi_forward_raw(f, b)
    once()
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(reps):
        once()
    e1.record()
    torch.cuda.synchronize()
    return e0.elapsed_time(e1) / reps


res = {}
for B in (1, 2, 4, 8):
    f, b = mk(B), mk(B)
    ops.set_segmentation(False)
    plain = step_ms(f, b)
    ops.set_segmentation(True)
    seg = step_ms(f, b)
    res[f"B{B}"] = {"sequences_per_launch": 2 * B * H, "ms_per_step_one_cta_per_sequence": plain, "ms_per_step": seg,
                    "frames_per_s": B * T / (seg * 1e-3), "speedup": plain / seg}
    del f, b
print(json.dumps({"workload": "configs[2] per-GPU share: B x T=30000, C=512, 8 heads, bf16, 12 blocks, both directions, fwd", **res}))
