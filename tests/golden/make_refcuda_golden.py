"""Generates tests/golden/refcuda_*.npz: outputs of the reference's OWN, UNMODIFIED CUDA kernels
(wenet/rwkv_v6/cuda/wkv6_cuda.cu, compiled by oracle/build_ref.py into oracle/_ref/libwkv6_ref.so) run on a B200.

    gpurun -- 'python tests/golden/make_refcuda_golden.py gpurun_out/golden'   # then copy the .npz files here

Inputs are NOT stored: they are regenerated from the seed by tests/_torch_ref.make_inputs (CPU generator,
deterministic), rounded once to the I/O dtype.  A checksum of the rounded inputs is stored so that a drift of the
generator is detected instead of silently comparing different data.  bf16 tensors are stored as their uint16 bits.
"""
import ctypes
import os
import sys

import numpy as np
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
from _torch_ref import make_inputs  # noqa: E402

CASES = [
    # name, dtype, B, T, H, w_mode, layer, seed, backward?, time stride stored
    ("f32_small", torch.float32, 1, 96, 2, "init", 3, 101, True, 1),
    ("bf16_small", torch.bfloat16, 2, 80, 2, "init", 5, 102, True, 1),
    ("f32_stress", torch.float32, 1, 64, 1, "stress", 0, 103, True, 1),
    ("bf16_stress", torch.bfloat16, 1, 70, 2, "stress", 11, 104, True, 1),
    ("f32_cfg1", torch.float32, 2, 250, 4, "init", 6, 105, False, 5),      # BASELINE configs[0] shape, fwd only
    ("bf16_cfg2row", torch.bfloat16, 1, 750, 8, "init", 9, 106, False, 10),  # one utterance of configs[1]
]


def rounded_inputs(dtype, B, T, H, w_mode, layer, seed):
    inp = make_inputs(B, T, H, 64, seed=seed, dtype=torch.float64, w_mode=w_mode, layer=layer)
    return {n: inp[n].to(dtype).contiguous() for n in ("r", "k", "v", "w", "u", "gy")}


def checksum(inp) -> float:
    return float(sum(inp[n].double().abs().sum() for n in ("r", "k", "v", "w", "u", "gy")))


def to_np(t: torch.Tensor) -> np.ndarray:
    t = t.detach().cpu().contiguous()
    return t.view(torch.int16).numpy().view(np.uint16) if t.dtype == torch.bfloat16 else t.numpy()


def main(out_dir: str) -> None:
    lib = ctypes.CDLL(os.path.join(ROOT, "oracle", "_ref", "libwkv6_ref.so"))
    os.makedirs(out_dir, exist_ok=True)
    p = lambda t: ctypes.c_void_p(t.data_ptr())
    for name, dtype, B, T, H, w_mode, layer, seed, bwd, stride in CASES:
        inp = rounded_inputs(dtype, B, T, H, w_mode, layer, seed)
        g = {n: t.cuda() for n, t in inp.items()}
        C = H * 64
        code = 0 if dtype == torch.bfloat16 else 1
        y = torch.empty_like(g["r"])
        torch.cuda.synchronize()
        rc = lib.wkv6_ref_forward(code, B, T, C, H, p(g["r"]), p(g["k"]), p(g["v"]), p(g["w"]), p(g["u"]), p(y))
        torch.cuda.synchronize()
        assert rc == 0, rc
        out = {"y": to_np(y[:, ::stride])}
        if bwd:
            gr, gk, gv, gw = (torch.empty_like(g["r"]) for _ in range(4))
            gu = torch.empty(B, C, device="cuda", dtype=dtype)
            rc = lib.wkv6_ref_backward(code, B, T, C, H, p(g["r"]), p(g["k"]), p(g["v"]), p(g["w"]), p(g["u"]),
                                       p(g["gy"]), p(gr), p(gk), p(gv), p(gw), p(gu))
            torch.cuda.synchronize()
            assert rc == 0, rc
            out.update(gr=to_np(gr), gk=to_np(gk), gv=to_np(gv), gw=to_np(gw), gu=to_np(gu))
        meta = np.array([B, T, H, layer, seed, stride, int(bwd)], dtype=np.int64)
        np.savez_compressed(os.path.join(out_dir, f"refcuda_{name}.npz"), meta=meta, w_mode=np.array(w_mode),
                            dtype=np.array(str(dtype)), checksum=np.array(checksum(inp)), **out)
        print("wrote", name, {k: v.shape for k, v in out.items()})


if __name__ == "__main__":
    main(sys.argv[1] if len(sys.argv) > 1 else os.path.join(ROOT, "gpurun_out", "golden"))
