"""Generates tests/golden/refpy_*.npz by running the reference's OWN Python modules, imported unmodified from
/root/reference in the build container (CPU): RWKV_Tmix_x060c inside RWKV_TmixWrapper, and the bidirectional /
direction-dropout wrappers.  The only thing the reference cannot do on a CPU is
the WKV6 op itself (CUDA only, model.py:105-106): here torch.ops.wkv6.* is backed by the CPU oracle, which is
pinned separately against the reference CUDA kernels (refcuda_*.npz).  Import shims: SURVEY.md App. B.

    python tests/golden/make_refpy_golden.py            # writes tests/golden/refpy_*.npz

Inputs and parameters are regenerated from seeds by the tests (the host mirror in
paper_accurate_fast_cheap_b200/wrappers.py draws its parameters in the same RNG order; this script asserts that).
"""
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
REF = "/root/reference"


def install_reference_shims():
    import oracle
    wenet = types.ModuleType("wenet")
    wenet.__path__ = [os.path.join(REF, "wenet")]
    sys.modules["wenet"] = wenet
    from torch.utils import cpp_extension
    cpp_extension.load = lambda *a, **k: None                       # model.py:105-106 JIT build -> no-op
    lib = torch.library.Library("wkv6", "DEF")
    sig_f = "(int B, int T, int C, int H, Tensor r, Tensor k, Tensor v, Tensor w, Tensor u, Tensor(a!) y) -> ()"
    sig_b = ("(int B, int T, int C, int H, Tensor r, Tensor k, Tensor v, Tensor w, Tensor u, Tensor gy, "
             "Tensor(a!) gr, Tensor(b!) gk, Tensor(c!) gv, Tensor(d!) gw, Tensor(e!) gu) -> ()")

    def fwd(B, T, C, H, r, k, v, w, u, y):
        out, _ = oracle.fwd(r, k, v, w, u, prec="f32")
        y.copy_(out.to(y.dtype))

    def bwd(B, T, C, H, r, k, v, w, u, gy, gr, gk, gv, gw, gu):
        outs = oracle.bwd_ref(r, k, v, w, u, gy, prec="f32")
        for dst, src in zip((gr, gk, gv, gw, gu), outs):
            dst.copy_(src.to(dst.dtype))

    for suffix in ("", "_fp32"):
        lib.define("forward" + suffix + sig_f)
        lib.define("backward" + suffix + sig_b)
        lib.impl("forward" + suffix, fwd, "CPU")
        lib.impl("backward" + suffix, bwd, "CPU")
    return lib


def ref_classes():
    from wenet.rwkv_v6.rwkv_wrapper import RWKV_TmixWrapper
    from wenet.rwkv_v6.rwkv_wrapper_bidirectional import RWKV_TmixWrapper_bidirectional
    from wenet.rwkv_v6.rwkv_wrapper_bidirectional_direction_dropout import \
        RWKV_TmixWrapper_bidirectional_direction_dropout
    from wenet.rwkv_v6.rwkv_wrapper_bidirectional_direction_dropout_both import \
        RWKV_TmixWrapper_bidirectional_direction_dropout_both
    return {"uni": RWKV_TmixWrapper, "bi": RWKV_TmixWrapper_bidirectional,
            "drop": RWKV_TmixWrapper_bidirectional_direction_dropout,
            "drop_both": RWKV_TmixWrapper_bidirectional_direction_dropout_both}


# name, class key, C, layer_id, do_bfloat16, B, T, train calls (0 = eval), stored time stride, env
CASES = [
    ("bi_f32_cfg1", "bi", 256, 2, False, 2, 250, 0, 5, {}),                 # BASELINE configs[0]: one bi block, CPU fp32
    ("bi_bf16", "bi", 128, 5, True, 2, 64, 0, 1, {}),
    ("uni_f32", "uni", 128, 0, False, 1, 70, 0, 1, {}),
    ("drop_eval_bi", "drop", 128, 7, True, 2, 50, 0, 1, {}),
    ("drop_eval_alt_r2l", "drop", 128, 3, False, 1, 40, 0, 1,
     {"RWKV_ALT_DECODING": "1", "RWKV_BIDIRECTIONAL_LAYERS": "0"}),
    ("drop_eval_l2r", "drop_both", 128, 3, False, 1, 40, 0, 1, {"RWKV_BIDIRECTIONAL_LAYERS": "0"}),
    ("drop_train", "drop", 128, 4, False, 1, 48, 6, 1, {}),
    ("drop_both_train", "drop_both", 128, 9, False, 1, 48, 10, 1, {}),
]
PARAM_SEED, X_SEED, CALL_SEED = 4242, 99, 5   # seed 5 draws 0,1,0,0,0,1,...: every branch is taken


def build(cls, C, layer_id, do_bf16):
    torch.manual_seed(PARAM_SEED + layer_id)
    return cls(64, C, 12, "v6", "bi", 2048, do_bf16, layer_id)


def main():
    keep_alive = install_reference_shims()   # the Library must outlive the calls
    import paper_accurate_fast_cheap_b200.wrappers as mine
    mine_cls = {"uni": mine.RWKV_TmixWrapper, "bi": mine.RWKV_TmixWrapper_bidirectional,
                "drop": mine.RWKV_TmixWrapper_bidirectional_direction_dropout,
                "drop_both": mine.RWKV_TmixWrapper_bidirectional_direction_dropout_both}
    refs = ref_classes()
    for name, key, C, layer_id, bf16, B, T, ncalls, stride, env in CASES:
        for k_ in ("RWKV_ALT_DECODING", "RWKV_BIDIRECTIONAL_LAYERS"):
            os.environ.pop(k_, None)
        os.environ.update(env)
        m = build(refs[key], C, layer_id, bf16)
        mm = build(mine_cls[key], C, layer_id, bf16)
        sd, sd2 = m.state_dict(), mm.state_dict()
        assert list(sd.keys()) == list(sd2.keys()), name
        for k_ in sd:
            assert torch.equal(sd[k_], sd2[k_]), (name, k_)
        x = torch.randn(B, T, C, generator=torch.Generator().manual_seed(X_SEED))
        outs = []
        if ncalls == 0:
            m.eval()
            with torch.no_grad():
                o, cache = m(x, x, x)
            outs.append(o)
        else:
            m.train()
            torch.manual_seed(CALL_SEED)
            with torch.no_grad():
                for _ in range(ncalls):
                    o, cache = m(x, x, x)
                    outs.append(o)
        out = torch.stack(outs, 0)[:, :, ::stride]
        assert out.dtype == torch.float32
        psum = float(sum(v.double().abs().sum() for v in sd.values()))
        np.savez_compressed(os.path.join(HERE, f"refpy_{name}.npz"), out=out.numpy(), key=np.array(key),
                            meta=np.array([C, layer_id, int(bf16), B, T, ncalls, stride], dtype=np.int64),
                            env=np.array(";".join(f"{a}={b}" for a, b in env.items())),
                            param_checksum=np.array(psum), x_checksum=np.array(float(x.double().abs().sum())))
        print("wrote", name, tuple(out.shape), "distinct outputs:", len({float(o.double().sum()) for o in outs}))


if __name__ == "__main__":
    main()
