"""Debug helper (GPU box): clock64() phase trace of CTA 0, chunk 4, of the chunk kernels on the bench workload."""
import ctypes, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if "WKV6_B200_LIB" not in os.environ:      # the in-loop stamps live in the trace build (paper_accurate_fast_cheap_b200/build.py --trace)
    _t = os.path.join(ROOT, "paper_accurate_fast_cheap_b200", "libwkv6_b200_trace.so")
    if os.path.exists(_t):
        os.environ["WKV6_B200_LIB"] = _t
    else:
        print("note: no trace build (python -m paper_accurate_fast_cheap_b200.build --trace): only the per-CTA timeline is recorded")
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
flush = torch.empty(512 << 20, dtype=torch.uint8, device="cuda")   # > 126 MB L2: the traced launches see cold inputs, like the bench
L.wkv6_b200_debug_trace(1, None)
flush.fill_(1)
pk.bi_forward_raw(f, r_, ckpt=ck); flush.fill_(2); pk.bi_backward_raw(f, r_, a["gy"], b["gy"], ckpt=ck)
torch.cuda.synchronize()
L.wkv6_b200_debug_trace(1, buf)
v = list(buf)
for name, base, n in (("fwd", 0, 15), ("bwd", 64, 16)):
    for who, off in (("t0 ", 0), ("t32", 16)):
        ts = v[base + off: base + off + n]
        t0 = ts[0]
        print(name, who, " ".join(f"{i}:{(x - t0) if x else -1}" for i, x in enumerate(ts)))
print("bwd loop tops of CTA 0 (cycles per chunk):", " ".join(str(v[96 + i + 1] - v[96 + i]) for i in range(31) if v[96 + i + 1] and v[96 + i]))
L.wkv6_b200_debug_trace(0, None)

# ---- per-CTA timeline of the forward chunk kernel (globaltimer ns): residency per SM, wave structure
L.wkv6_b200_debug_trace(1, None)
flush.fill_(3)
pk.bi_forward_raw(f, r_)
torch.cuda.synchronize()
cb = (ctypes.c_ulonglong * 4096)()
L.wkv6_b200_debug_trace_ctas(cb)
L.wkv6_b200_debug_trace(1, buf)
v = list(buf)
print("fwd loop tops (cycles since chunk 0):", " ".join(str(x - v[32]) for x in v[32:32 + 12]))
L.wkv6_b200_debug_trace(0, None)
rows = [(cb[4 * i], cb[4 * i + 1], cb[4 * i + 2]) for i in range(2 * B * H) if cb[4 * i]]
if rows:
    t_min = min(r[0] for r in rows); t_max = max(r[2] for r in rows)
    durs = sorted(r[2] - r[0] for r in rows)
    starts = sorted(r[0] - t_min for r in rows)
    per_sm = {}
    for s, sm, e in rows:
        per_sm.setdefault(sm, []).append((s - t_min, e - t_min))
    overlap2 = sum(1 for v in per_sm.values() if len(v) >= 2 and sorted(v)[1][0] < sorted(v)[0][1])
    solo = sorted(v[0][1] - v[0][0] for v in per_sm.values() if len(v) == 1)
    pair = sorted(e - s for v in per_sm.values() if len(v) == 2 for s, e in v)
    if solo and pair:
        print(f"solo CTA ns min/med/max {solo[0]}/{solo[len(solo)//2]}/{solo[-1]}; paired CTA ns {pair[0]}/{pair[len(pair)//2]}/{pair[-1]}; cta0 {rows[0][2]-rows[0][0]}")
    full = [(cb[4 * i + 2] - cb[4 * i], i, cb[4 * i + 1], cb[4 * i] - t_min) for i in range(2 * B * H) if cb[4 * i]]
    by_sm = {}
    for d_, i, sm_, st_ in full:
        by_sm.setdefault(sm_, []).append((i, d_))
    print("slowest CTAs (dur ns, block, sm, start, co-resident blocks):")
    for d_, i, sm_, st_ in sorted(full, reverse=True)[:8]:
        print("  ", d_, i, sm_, st_, [x for x in by_sm[sm_] if x[0] != i])
    hist = {}
    for d_, *_ in full:
        hist[d_ // 4000 * 4] = hist.get(d_ // 4000 * 4, 0) + 1
    print("duration histogram (us bucket: count):", sorted(hist.items()))
    print(f"ctas {len(rows)} span {t_max - t_min} ns; cta duration min/med/max {durs[0]}/{durs[len(durs)//2]}/{durs[-1]} ns; "
          f"SMs used {len(per_sm)}, SMs with 2 CTAs {sum(1 for v in per_sm.values() if len(v) >= 2)}, overlapping in time {overlap2}; "
          f"last start {starts[-1]} ns")


# ---- per-CTA timeline of the backward chunk kernel (persistent balanced grid)
pk.bi_forward_raw(f, r_, ckpt=ck)
flush.fill_(4)
torch.cuda.synchronize()
L.wkv6_b200_debug_trace(0, None)
L.wkv6_b200_debug_trace(1, None)
pk.bi_backward_raw(f, r_, a["gy"], b["gy"], ckpt=ck)
torch.cuda.synchronize()
L.wkv6_b200_debug_trace_ctas(cb)
rows = [(cb[4 * i], cb[4 * i + 1], cb[4 * i + 2], cb[4 * i + 3]) for i in range(1024) if cb[4 * i] and cb[4 * i + 2]]
if rows:
    t_min = min(r[0] for r in rows); t_max = max(r[2] for r in rows)
    durs = sorted(r[2] - r[0] for r in rows)
    print(f"bwd: {len(rows)} CTAs on {len(set(r[1] for r in rows))} SMs, kernel span {t_max - t_min} ns; CTA ns min/med/max "
          f"{durs[0]}/{durs[len(durs)//2]}/{durs[-1]}; mean {sum(durs)//len(durs)}; start spread {max(r[0] for r in rows) - t_min} ns; "
          f"end spread {t_max - min(r[2] for r in rows)} ns; hand-over wait ns max {max(r[3] for r in rows)} "
          f"sum {sum(r[3] for r in rows)}")
    full = [(cb[4 * i + 2] - cb[4 * i], i, cb[4 * i + 1]) for i in range(1024) if cb[4 * i] and cb[4 * i + 2]]
    print("bwd per CTA (blockIdx:smid:ns):", " ".join(f"{i}:{sm_}:{d_}" for d_, i, sm_ in sorted(full, key=lambda x: x[1])))
L.wkv6_b200_debug_trace(0, None)
