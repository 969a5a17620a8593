"""Times the reference's own CUDA WKV6 kernels (oracle/_ref/libwkv6_ref.so, built by oracle/build_ref.py from
/root/reference/wenet/rwkv_v6/cuda/wkv6_cuda.cu) on this GPU for the bench workload (configs[1]: B=16, T=750,
C=512, H=8, bf16), one direction per launch as the reference does, plus the flip/add/cast the reference wrapper
adds per layer (rwkv_wrapper_bidirectional.py:43-56).  Measurement helper (not product, not oracle-as-product):
prints one JSON line.  Usage on the box:  python scripts/time_ref_cuda.py"""
import ctypes
import json
import os
import sys

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)


def main():
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libwkv6_ref.so"))
    B, T, C, H = 16, 750, 512, 8
    dev = "cuda:0"
    g = torch.Generator(device="cpu").manual_seed(1)
    mk = lambda: torch.randn(B, T, C, generator=g).to(torch.bfloat16).to(dev)
    r, k, v, gy = mk(), mk(), mk(), mk()
    w = (-3 + 0.5 * torch.randn(B, T, C, generator=g)).to(torch.bfloat16).to(dev)
    u = torch.randn(H, 64, generator=g).to(torch.bfloat16).to(dev)
    y = torch.empty_like(r)
    outs = [torch.empty_like(r) for _ in range(4)] + [torch.empty(B, C, device=dev, dtype=torch.bfloat16)]
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)

    def fwd():
        lib.wkv6_ref_forward(0, B, T, C, H, p(r), p(k), p(v), p(w), p(u), p(y))

    def bwd():
        lib.wkv6_ref_backward(0, B, T, C, H, p(r), p(k), p(v), p(w), p(u), p(gy), *[p(o) for o in outs])

    def wrapper_overhead():
        xf = torch.flip(r, [1])
        yb = torch.flip(y, [1])
        return ((y + yb) / 2).float(), xf

    def timeit(fn, n=10):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        ts = []
        for _ in range(n):
            flush.zero_()
            a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            a.record(); fn(); b.record()
            torch.cuda.synchronize()
            ts.append(a.elapsed_time(b))
        ts.sort()
        return ts[len(ts) // 2]

    with torch.cuda.stream(torch.cuda.default_stream()):
        t_f, t_b, t_w = timeit(fwd), timeit(bwd), timeit(wrapper_overhead)
    frames = B * T
    line = {"what": "reference CUDA WKV6 kernels on this GPU, one direction, bf16, B=16 T=750 C=512 H=8",
            "gpu": torch.cuda.get_device_name(0), "fwd_ms": t_f, "bwd_ms": t_b, "wrapper_flip_add_cast_ms": t_w,
            "fwd_gbs": 5 * frames * C * 2 / t_f / 1e6, "bwd_gbs": 9 * frames * C * 2 / t_b / 1e6,
            "bi_fwd_frames_per_s_12_blocks": frames / (12 * (2 * t_f + t_w) * 1e-3),
            "bi_fwdbwd_frames_per_s_12_blocks": frames / (12 * (2 * (t_f + t_b) + 2 * t_w) * 1e-3)}
    print(json.dumps(line))


if __name__ == "__main__":
    main()
