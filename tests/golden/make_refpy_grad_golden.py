"""Generates tests/golden/refpygrad_*.npz: INPUT and PARAMETER GRADIENTS of the reference's own Python modules
(bidirectional wrapper and both direction-dropout wrappers, imported unmodified from /root/reference in the build
container) after `(out * gy).sum().backward()`, training mode, including steps where one direction is dropped and
its 17 parameters come back with `grad is None` (rwkv_wrapper_bidirectional_direction_dropout.py:59-69,
..._both.py:53-71; autograd through model.py:136-152).  torch.ops.wkv6.{forward,backward}* are backed by the CPU
oracle's literal restatement of the reference kernels (shims: make_refpy_golden.py).

    python tests/golden/make_refpy_grad_golden.py

Stored per case: for the FIRST step of every direction outcome (both / L2R only / R2L only) the output, x.grad and
every parameter gradient (tensors of >= 4096 elements flattened and sampled with stride 7), plus the per-step outcome
sequence.  Inputs and parameters are regenerated from seeds by the tests.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)
from make_refpy_golden import CALL_SEED, X_SEED, build, install_reference_shims, ref_classes  # noqa: E402

GY_SEED = 123
BIG, STRIDE = 4096, 7

# name, class key, C, layer_id, do_bfloat16, B, T, training calls
CASES = [
    ("bi_f32", "bi", 128, 2, False, 2, 40, 1),
    ("bi_bf16", "bi", 128, 5, True, 2, 64, 1),
    ("drop_f32", "drop", 128, 4, False, 1, 48, 6),
    ("drop_both_f32", "drop_both", 128, 9, False, 1, 48, 10),
    ("drop_both_bf16", "drop_both", 128, 9, True, 2, 48, 10),
]


def sample(t):
    t = t.detach().float().reshape(-1)
    return (t[::STRIDE] if t.numel() >= BIG else t).numpy()


def main():
    keep_alive = install_reference_shims()   # noqa: F841  (the Library must outlive the calls)
    refs = ref_classes()
    for name, key, C, layer_id, bf16, B, T, ncalls in CASES:
        for k_ in ("RWKV_ALT_DECODING", "RWKV_BIDIRECTIONAL_LAYERS"):
            os.environ.pop(k_, None)
        m = build(refs[key], C, layer_id, bf16)
        m.train()
        x0 = torch.randn(B, T, C, generator=torch.Generator().manual_seed(X_SEED))
        gy = torch.randn(B, T, C, generator=torch.Generator().manual_seed(GY_SEED))
        torch.manual_seed(CALL_SEED)
        names = [n for n, _ in m.named_parameters()]
        store, outcomes = {}, []
        for step in range(ncalls):
            m.zero_grad(set_to_none=True)
            x = x0.clone().requires_grad_(True)
            out, _ = m(x, x, x)
            (out.float() * gy).sum().backward()
            has_f = any(p.grad is not None for n, p in m.named_parameters() if n.startswith("rwkv_wrapper_forward"))
            has_b = any(p.grad is not None for n, p in m.named_parameters() if n.startswith("rwkv_wrapper_backward"))
            oc = (1 if has_f else 0) | (2 if has_b else 0)
            outcomes.append(oc)
            if f"s{oc}_step" in store:
                continue
            store[f"s{oc}_step"] = np.array(step)
            store[f"s{oc}_out"] = out.detach().float().numpy()
            store[f"s{oc}_gx"] = x.grad.float().numpy()
            for i, (n, p) in enumerate(m.named_parameters()):
                if p.grad is not None:
                    store[f"s{oc}_g{i}"] = sample(p.grad)
        np.savez_compressed(os.path.join(HERE, f"refpygrad_{name}.npz"), key=np.array(key), names=np.array(names),
                            meta=np.array([C, layer_id, int(bf16), B, T, ncalls], dtype=np.int64),
                            outcomes=np.array(outcomes, dtype=np.int64), **store)
        print("wrote", name, "outcomes", outcomes, "stored", sorted(k for k in store if k.endswith("_step")))


if __name__ == "__main__":
    main()
