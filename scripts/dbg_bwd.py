"""Debug helper (GPU box): per-tensor errors of the bf16 backward (chunk vs scan kernels) against the fp64 oracle."""
import sys, os
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import torch
import oracle
import paper_accurate_fast_cheap_b200 as pk
from paper_accurate_fast_cheap_b200 import _lib
from test_gpu_parity import _inputs, _cuda

cases = [(1, 64, 1, False, "init"), (1, 16, 1, False, "init"), (2, 250, 4, False, "init"), (2, 130, 2, True, "init"),
         (3, 53, 2, True, "stress"), (2, 750, 8, False, "init"), (1, 1, 1, False, "init")]
if len(sys.argv) > 1:
    cases = cases[:int(sys.argv[1])]
for (B, T, H, st, wm) in cases:
    for rev in (False, True):
        inp = _inputs(B, T, H, torch.bfloat16, wm, st)
        g = _cuda(inp)
        ref = oracle.bwd(inp["r"], inp["k"], inp["v"], inp["w"], inp["u"], inp["gy"], s0=inp.get("s0"), gsT=inp.get("gsT"),
                         rev=rev, prec="f64")
        for algo, name in ((_lib.ALGO_SCAN, "scan "), (_lib.ALGO_AUTO, "chunk")):
            pk.set_algo(algo)
            out = pk.wkv6_backward_raw(g["r"], g["k"], g["v"], g["w"], g["u"], g["gy"], s0=g.get("s0"), gsT=g.get("gsT"),
                                       reverse=rev, want_gs0=st)
            torch.cuda.synchronize()
            errs = {n: oracle.rel_err(a, b) for n, a, b in zip(("gr", "gk", "gv", "gw", "gu"), out[:5], ref[:5])}
            if st:
                errs["gs0"] = oracle.rel_err(out[5], ref[5])
            bad = any((e != e) or e > 2e-2 for e in errs.values())
            print(f"B{B} T{T} H{H} st{int(st)} {wm:6s} rev{int(rev)} {name}", " ".join(f"{n}={e:.1e}" for n, e in errs.items()),
                  "  <-- BAD" if bad else "", flush=True)
        pk.set_algo(_lib.ALGO_AUTO)
