#!/usr/bin/env python
"""bench.py -- encoder frames/s of the bidirectional WKV6 hot path on N B200s (one process per GPU).

Workload (BASELINE.json configs[1]): full bidirectional RWKV-6 encoder's WKV6 work -- 12 blocks, d_model 512,
8 heads x 64, bf16, B=16 x T=750 subsampled frames (30 s), inference (forward, both directions, one launch per
block).  One "step" = the 12 bidirectional scans of one batch.  N>1: every rank scans its own batch of 16
utterances (sharded by utterance, no collective) -> weak scaling.  `--mode train` times fwd+bwd instead.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--mode infer|train]
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

CFG = dict(layers=12, C=512, H=8, B=16, T=750)


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=200)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--mode", default="infer", choices=["infer", "train"])
    ap.add_argument("--algo", default="auto", choices=["auto", "scan", "chunk"])
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-train", action="store_true", help="skip the fwd+bwd figures reported beside the inference line")
    ap.add_argument("--no-longform", action="store_true", help="skip the configs[2] (T = 30 000) figure reported beside it")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons with nvidia-smi while the timed region runs."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown," \
        "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap"

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.rows, self.stop_flag = index, [], False

    def run(self):
        while not self.stop_flag:
            try:
                out = subprocess.run(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                      "-i", str(self.index)], capture_output=True, text=True, timeout=5).stdout
                self.rows.append([x.strip() for x in out.strip().split(",")])
            except Exception:
                pass
            time.sleep(0.05)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


_CPU_SETS = None


def cpu_sample(passes=1, train=False):
    """The oracle port (C restatement of the reference kernels, OpenMP over batch x head, all host cores) timed on
    `passes` full passes of the workload: 12 blocks x both directions x full B x T, fp32.  Returns frames/s."""
    global _CPU_SETS
    import torch
    import oracle
    from _torch_ref import make_inputs
    c = CFG
    if _CPU_SETS is None:
        _CPU_SETS = [make_inputs(c["B"], c["T"], c["H"], 64, seed=s, dtype=torch.float32, layer=s % 12) for s in (11, 12)]
        s0 = _CPU_SETS[0]
        oracle.fwd(s0["r"][:1], s0["k"][:1], s0["v"][:1], s0["w"][:1], s0["u"])   # warm (builds / loads the library)
    t0 = time.perf_counter()
    for _ in range(passes):
        for layer in range(c["layers"]):
            for s, rev in zip(_CPU_SETS, (False, True)):
                oracle.fwd(s["r"], s["k"], s["v"], s["w"], s["u"], rev=rev, prec="f32")
                if train:
                    oracle.bwd(s["r"], s["k"], s["v"], s["w"], s["u"], s["gy"], rev=rev, prec="f32")
    dt = time.perf_counter() - t0
    return passes * c["B"] * c["T"] / dt, dt, oracle.num_threads()


def cpu_baseline_bounded(budget_s=10.0, train=False):
    """About `budget_s` seconds of CPU work: as many full passes as fit."""
    fps1, dt1, th = cpu_sample(1, train)
    n = max(1, min(200, int(budget_s / max(dt1, 1e-3))))
    fps, dt, th = cpu_sample(n, train)
    return fps, dt, th, n


def config_dict(args, extra=None):
    c = CFG
    d = {"workload": f"configs[1]: bi-WKV6 of a {c['layers']}-block RWKV-6 encoder, d_model={c['C']}, {c['H']} heads x 64, "
                     f"B={c['B']} x T={c['T']} per GPU, {'fwd (inference)' if args.mode == 'infer' else 'fwd+bwd (training)'}",
         "layers": c["layers"], "B_per_gpu": c["B"], "T": c["T"], "C": c["C"], "H": c["H"], "mode": args.mode,
         "parallelism": f"utterance-sharded x{args.gpus}, no collective",
         "l2": "inputs (1.18 GB per step) exceed the 126 MB L2; every block reads its own tensors"}
    if extra:
        d.update(extra)
    return d


_OUT_FD = None


def claim_stdout():
    """stdout carries exactly ONE JSON line: everything else a library may print there (NCCL's version banner, build
    chatter) is sent to stderr for the whole run; emit() writes the line to the real stdout."""
    global _OUT_FD
    sys.stdout.flush()
    _OUT_FD = os.dup(1)
    os.dup2(2, 1)


def emit(line):
    sys.stdout.flush()
    data = (json.dumps(line) + "\n").encode()
    if _OUT_FD is None:
        os.write(1, data)
    else:
        os.write(_OUT_FD, data)


def run_reference(args):
    """Reference arm: the reference has no CPU WKV6 (model.py:105-106 is CUDA-only), so this times the oracle
    port of its kernels on all host cores.  Rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    import oracle
    oracle.build()
    # torchrun exports OMP_NUM_THREADS=1 to its workers: the reference arm uses every host core it can, as at N=1
    oracle.set_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
    train = args.mode == "train"
    steps = min(args.steps, 20)          # one step = one full pass on the CPU (~0.25 s fwd); bounded to a few minutes
    warm = min(args.warmup, 2)
    vals = []
    for i in range(warm + steps):
        fps, dt, th = cpu_sample(1, train)
        if i >= warm:
            vals.append((fps, dt))
    tot = sum(v[1] for v in vals)
    fps = len(vals) * CFG["B"] * CFG["T"] / tot
    ms = 1e3 * tot / len(vals)
    sample = (f"{len(vals)} full passes (12 blocks x both directions x B=16 x T=750), fp32 oracle port of the reference "
              "kernels, OpenMP over batch x head on all host cores (the reference itself has no CPU WKV6)")
    line = {"impl": "reference", "metric": "encoder frames/s (bi-WKV6)", "value": fps, "unit": "frames/s",
            "n_gpus": args.gpus, "steps": steps, "warmup": warm, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": th, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


def make_layer_inputs(torch, dev, layer, dtype, train, pinned=False):
    """Synthetic op-level inputs of SURVEY 8(d) for one block: L2R set and R2L set."""
    c = CFG
    g = torch.Generator(device="cpu").manual_seed(777 + layer)
    C, H, B, T, L = c["C"], c["H"], c["B"], c["T"], c["layers"]
    ratio = layer / (L - 1)
    n = torch.arange(C, dtype=torch.float32)
    base = (-6 + 5 * (n / (C - 1)) ** (0.7 + 1.3 * ratio)).view(1, 1, C)
    u = (ratio * (1 - n / (C - 1)) + ((n + 1) % 3 - 1) * 0.1).view(H, 64)
    sets = []
    for _ in range(2):
        t = [torch.randn(B, T, C, generator=g) for _ in range(3)]
        t.append(base + 0.5 * torch.randn(B, T, C, generator=g))
        t.append(u.clone())
        if train:
            t.append(torch.randn(B, T, C, generator=g))
        t = [x.to(dtype).contiguous() for x in t]
        sets.append([x.pin_memory() for x in t] if pinned else [x.to(dev) for x in t])
    return sets


def run_b200(args):
    import torch
    import torch.distributed as dist
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import _lib

    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    pk.set_algo({"auto": _lib.ALGO_AUTO, "scan": _lib.ALGO_SCAN, "chunk": _lib.ALGO_CHUNK}[args.algo])
    L = _lib.lib()
    c = CFG
    B, T, C, H, NL = c["B"], c["T"], c["C"], c["H"], c["layers"]
    dt = torch.bfloat16
    layers = [make_layer_inputs(torch, dev, l, dt, True) for l in range(NL)]      # gy is only read by the backward
    mk = lambda: torch.empty(B, T, C, device=dev, dtype=dt)
    outs = [[mk(), mk()] for _ in range(NL)]
    gouts = [[[mk() for _ in range(4)] + [torch.empty(B, C, device=dev, dtype=dt)] for _ in range(2)] for _ in range(NL)]
    stream = torch.cuda.current_stream().cuda_stream
    flags = _lib.BI_L2R | _lib.BI_R2L
    p = lambda t: t.data_ptr()
    ck_bytes = L.wkv6_b200_checkpoint_bytes(B, T, C, H)
    ckpts = [torch.empty((2, max(ck_bytes, 16)), device=dev, dtype=torch.uint8) for _ in range(NL)]
    ws_bytes = L.wkv6_b200_backward_workspace_bytes(0, B, T, C, H, 2)
    ws = torch.empty(max(ws_bytes, 16), device=dev, dtype=torch.uint8)

    def fwd_layer(l, train, ins=None, ys=None):
        f, b = ins or layers[l]
        yf, yb = ys or outs[l]
        ckf = p(ckpts[l][0]) if train else None
        ckb = p(ckpts[l][1]) if train else None
        _lib.check(L.wkv6_b200_bi_forward(0, B, T, C, H, p(f[0]), p(f[1]), p(f[2]), p(f[3]), p(f[4]), None, p(yf), None, ckf,
                                          p(b[0]), p(b[1]), p(b[2]), p(b[3]), p(b[4]), None, p(yb), None, ckb, flags, stream))

    def bwd_layer(l):
        f, b = layers[l]
        gf, gb = gouts[l]
        _lib.check(L.wkv6_b200_bi_backward(
            0, B, T, C, H,
            p(f[0]), p(f[1]), p(f[2]), p(f[3]), p(f[4]), None, p(f[5]), None, p(ckpts[l][0]), *[p(x) for x in gf], None,
            p(b[0]), p(b[1]), p(b[2]), p(b[3]), p(b[4]), None, p(b[5]), None, p(ckpts[l][1]), *[p(x) for x in gb], None,
            flags, p(ws), ws_bytes, stream))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(ms):
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        if world > 1:
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def timed(train, steps, warmup, sampler=None):
        """`steps` steps of the 12-block workload (fwd, or fwd then bwd in reverse block order), the timed region
        bracketed by CUDA events on the launching stream with nothing but the kernel launches inside (an event record
        between two launches serialises them and costs ~2 us each).  Kernel durations for the roofline: the inference
        step is NL identical forward launches, so the average launch duration IS region time / launches; the training
        step mixes two kernels, so right after the timed region the same launches are timed again as two homogeneous
        regions (all forward launches of a step, then all backward launches), `steps` times each.
        Returns ms/step (max over ranks), avg fwd-launch ms, avg bwd-launch ms, launches in the timed region."""
        def step():
            for l in range(NL):
                fwd_layer(l, train)
            if train:
                for l in reversed(range(NL)):
                    bwd_layer(l)

        def region(fn, reps):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            torch.cuda.synchronize()
            e0.record()
            for _ in range(reps):
                fn()
            e1.record()
            torch.cuda.synchronize()
            return e0.elapsed_time(e1) / reps

        for _ in range(warmup):
            step()
        barrier()
        n0 = pk.launch_count()
        if sampler:
            sampler.start()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        t0.record()
        for _ in range(steps):
            step()
        t1.record()
        barrier()
        launches = pk.launch_count() - n0
        local_ms = t0.elapsed_time(t1) / steps
        ms = allmax(t0.elapsed_time(t1)) / steps
        if not train:
            return ms, local_ms / NL, None, launches
        fk = region(lambda: [fwd_layer(l, True) for l in range(NL)], steps) / NL
        bk = region(lambda: [bwd_layer(l) for l in reversed(range(NL))], steps) / NL
        return ms, fk, bk, launches

    train = args.mode == "train"
    sampler = ClockSampler(local)
    ms_step, fwd_ms, bwd_ms, launches = timed(train, args.steps, args.warmup, sampler)
    if len(sampler.rows) < 3:            # a short timed region: keep the same load running until the sampler has seen it
        t_end = time.time() + 1.0
        while time.time() < t_end and len(sampler.rows) < 4:
            for l in range(NL):
                fwd_layer(l, train)
            torch.cuda.synchronize()
    sampler.stop_flag = True
    value = world * B * T / (ms_step * 1e-3)
    pk_, how = peaks()
    fwd_bytes = 2 * 5 * B * T * C * 2                     # both directions: read r,k,v,w + write y, bf16
    bwd_bytes = 2 * 9 * B * T * C * 2                     # read r,k,v,w,gy + write gr,gk,gv,gw

    # dram__bytes_read.sum + dram__bytes_write.sum per launch from the ncu --set full capture of this workload
    # (profiles/r1_chunk_kernels_ncu_summary.txt): fwd 98.5 + 19.0 MB, bwd 148.3 + 62.4 MB (outputs partly still in L2)
    NCU_TRAFFIC = {"fwd": 117.49e6, "bwd": 210.71e6}

    def roof(kernel, nbytes, ms, traffic=None):
        ach = nbytes / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": pk_["hbm_gbs"], "unit": "GB/s",
                "frac": ach / pk_["hbm_gbs"], "peak_source": how, "traffic": traffic, "alg_bytes_per_launch": nbytes,
                "kernel_ms": ms}

    roofline = roof("wkv6_fwd_chunk_lean_kernel: bidirectional forward, one launch per block", fwd_bytes, fwd_ms,
                    NCU_TRAFFIC["fwd"])
    extra = {}
    if train:
        roofline = roof("wkv6_bwd_chunk_kernel: bidirectional backward, one launch per block", bwd_bytes, bwd_ms,
                        NCU_TRAFFIC["bwd"])
        extra["roofline_fwd"] = roof("wkv6_fwd_chunk_lean_kernel (+ checkpoint stores)", fwd_bytes, fwd_ms)
        extra["roofline_fwd_bwd"] = roof("forward + backward launches of one block", fwd_bytes + bwd_bytes, fwd_ms + bwd_ms)
    elif not args.no_train:
        # the same workload as a training step (fwd + bwd), reported beside the inference headline
        tsteps = max(3, min(args.steps, 30))
        tms, tf, tb, _ = timed(True, tsteps, 3)
        extra["train"] = {"metric": "encoder frames/s (bi-WKV6 fwd+bwd)", "value": world * B * T / (tms * 1e-3),
                          "unit": "frames/s", "ms_per_step": tms, "steps": tsteps,
                          "roofline_bwd": roof("wkv6_bwd_chunk_kernel", bwd_bytes, tb, NCU_TRAFFIC["bwd"]),
                          "roofline_fwd_bwd": roof("forward + backward launches of one block", fwd_bytes + bwd_bytes, tf + tb)}

    # ---- long-form (BASELINE configs[2]): 8 utterances of T = 30 000 frames in total, sharded by utterance over the
    # ranks (8 / N per GPU: strong scaling), 12 blocks, both directions, forward; one input set per rank (2.4 GB per
    # launch at B = 8: nothing stays in the 126 MB L2).  Few long sequences per GPU: the library runs them time-segmented
    # on the scratch buffer lent here (wkv6_b200_set_scratch).
    if not train and not args.no_longform and world <= 8:
        LT, LB = 30000, max(1, 8 // world)
        gl = torch.Generator(device=dev).manual_seed(777 + rank)
        lsets = []
        for _ in range(2):
            t = [torch.randn(LB, LT, C, device=dev, generator=gl).to(dt) for _ in range(3)]
            t.append((-6 + 5 * torch.rand(LB, LT, C, device=dev, generator=gl)).to(dt))
            t.append((0.3 * torch.randn(H, 64, device=dev, generator=gl)).to(dt))
            lsets.append(t)
        ly = [torch.empty(LB, LT, C, device=dev, dtype=dt) for _ in range(2)]
        scratch = torch.empty(6 << 20, device=dev, dtype=torch.uint8)
        _lib.check(L.wkv6_b200_set_scratch(p(scratch), scratch.numel()))

        def lstep():
            f, b = lsets
            for _ in range(NL):
                _lib.check(L.wkv6_b200_bi_forward(0, LB, LT, C, H, p(f[0]), p(f[1]), p(f[2]), p(f[3]), p(f[4]), None, p(ly[0]), None,
                                                  None, p(b[0]), p(b[1]), p(b[2]), p(b[3]), p(b[4]), None, p(ly[1]), None, None,
                                                  flags, stream))

        lstep(); lstep(); barrier()
        n0 = pk.launch_count()
        a0, a1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        lreps = 3
        a0.record()
        for _ in range(lreps):
            lstep()
        a1.record(); barrier()
        lms = allmax(a0.elapsed_time(a1)) / lreps
        extra["longform"] = {"metric": "encoder frames/s (bi-WKV6)", "value": world * LB * LT / (lms * 1e-3), "unit": "frames/s",
                             "workload": f"configs[2]: 8 utterances x T=30000 in total, {LB} per GPU, C=512, 8 heads, bf16, 12 blocks, fwd",
                             "scaling": "strong", "ms_per_step": lms, "steps": lreps,
                             "launches_per_block": (pk.launch_count() - n0) // (lreps * NL)}
        _lib.check(L.wkv6_b200_set_scratch(None, 0))
        del lsets, ly, scratch

    # ---- end to end: pinned host inputs -> H2D -> op -> D2H of y, every step, through the C ABI
    e2e = None
    if not args.no_e2e and not train:
        host = [make_layer_inputs(torch, dev, l, dt, False, pinned=True) for l in range(2)]   # 2 distinct blocks, reused
        stage = [[[torch.empty_like(x, device=dev) for x in s_] for s_ in host[0]] for _ in range(2)]
        ystage = [[mk(), mk()] for _ in range(2)]
        yhost = [[torch.empty(B, T, C, dtype=dt).pin_memory() for _ in range(2)] for _ in range(2)]
        h2d = sum(x.numel() * x.element_size() for s_ in host[0] for x in s_) * NL
        d2h = 2 * B * T * C * 2 * NL

        # three streams, double-buffered staging: H2D of block l+1 and D2H of block l-1 overlap the scan of block l
        # (PCIe is full duplex); events order copy -> scan -> copy-back and guard the reuse of each staging slot
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        s_cmp = torch.cuda.current_stream()
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_cmp = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]

        def e2e_step():
            for l in range(NL):
                k = l & 1
                src, dst = host[k], stage[k]
                with torch.cuda.stream(s_in):
                    s_in.wait_event(ev_cmp[k])            # the scan that last read this staging slot is done
                    for s_h, s_d in zip(src, dst):
                        for xh, xd in zip(s_h, s_d):
                            xd.copy_(xh, non_blocking=True)
                    ev_in[k].record(s_in)
                s_cmp.wait_event(ev_in[k])
                s_cmp.wait_event(ev_out[k])               # y staging slot k has been copied back
                fwd_layer(l, False, ins=dst, ys=ystage[k])
                ev_cmp[k].record(s_cmp)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_cmp[k])
                    for yh, yd in zip(yhost[k], ystage[k]):
                        yh.copy_(yd, non_blocking=True)
                    ev_out[k].record(s_out)
            s_cmp.wait_stream(s_out)                      # the step ends when its last result is on the host

        e2e_steps = max(2, min(args.steps, 5))
        e2e_step(); barrier()
        a = torch.cuda.Event(enable_timing=True); b_ = torch.cuda.Event(enable_timing=True)
        a.record()
        for _ in range(e2e_steps):
            e2e_step()
        b_.record(); barrier()
        te = allmax(a.elapsed_time(b_))
        e2e = {"value": world * B * T / (te / e2e_steps * 1e-3), "unit": "frames/s",
               "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h, "steps": e2e_steps}

    if rank == 0:
        cpu = None
        if world == 1 and not args.no_cpu_baseline:
            import oracle
            oracle.set_threads(len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1))
            fps, secs, th, n = cpu_baseline_bounded(10.0, train)
            cpu = {"value": fps, "unit": "frames/s", "cores": th, "kind": "port",
                   "sample": f"{n} full passes of the workload (12 blocks x both directions x B=16 x T=750, fp32"
                             f"{', fwd+bwd' if train else ', fwd'}): {secs:.1f} s of CPU work on {th} threads"}
        line = {"metric": "encoder frames/s (bi-WKV6%s)" % (" fwd+bwd" if train else ""), "value": value, "unit": "frames/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": config_dict(args, {"algo": args.algo}), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
                "gpu_launches": launches, "clocks": sampler.summary()}
        line.update(extra)
        emit(line)
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    a = parse()
    claim_stdout()
    if a.impl == "reference":
        run_reference(a)
    else:
        run_b200(a)
