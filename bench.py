#!/usr/bin/env python
"""bench.py -- encoder frames/s of the bidirectional WKV6 hot path on N B200s (one process per GPU).

Headline (BASELINE.json `metric`: "encoder frames/s (bi-WKV6 fwd+bwd)"), measured on the shape of configs[1]: the
WKV6 work of a 12-block bidirectional RWKV-6 encoder, d_model 512, 8 heads x 64, bf16, B=16 x T=750 subsampled frames
per GPU -- per block ONE forward launch (both directions, chunk-state checkpoints stored) and ONE backward launch.
One "step" = the 12 forward launches followed by the 12 backward launches of one batch.  N>1: every rank scans its own
batch (sharded by utterance, no data-path collective) -> weak scaling.

Beside the headline the same JSON line carries (each is a leg of `--legs`, all on by default):
  infer     the forward alone (inference; configs[1] literally)
  e2e       the headline step through the C ABI with HOST buffers: pinned slab -> H2D -> fwd -> bwd -> D2H, every step
  cfg4      configs[3]: B=8 per GPU, direction dropout (seeded CPU Bernoulli sequence of ..._direction_dropout_both.py)
            as the launches' flags, and the NCCL gradient all-reduce of the 24 x 1 249 792 bf16 Tmix gradients in
            DDP-sized buckets (4 blocks = 20 MB), each issued on a side stream as soon as the backward launch of its last
            block is queued (DDP's find_unused_parameters=True volume: a dropped direction contributes zeros)
  cfg5      configs[4]: fp32, B=1, T=90 000, d_model 1024, 16 heads, one bidirectional block, forward
  longform  configs[2]: 8 utterances x T=30 000 in total, sharded over the ranks (strong scaling), forward
  conv      SURVEY 8f4: the Conformer ConvolutionModule that follows the time-mix in every layer (kernel 31, layer_norm,
            swish) on configs[1]'s shape, fp32: the fused GLU + depthwise conv + LayerNorm + swish core against its
            roofline, and the whole module (with its two pointwise GEMMs) against the reference's eager op sequence
  module    the attention-slot MODULES (the call wenet's encoder layer makes): 12 x RWKV_TmixWrapper_bidirectional, inference,
            and 12 x ..._direction_dropout_both, training step, bf16, B=16 (resp. 8) x T=750 -- this package against the
            reference's formulation on the same GPU (eager prologue / epilogue ops, its CUDA kernels, flips, add, cast)
  ref_cuda  the reference's own CUDA kernels (oracle/_ref/libwkv6_ref.so = wkv6_cuda.cu unmodified) on the same GPU
  cpu       the CPU port of the reference kernels (oracle/) on the host cores, bounded sample

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference] [--legs a,b,...]
"""
import argparse
import ctypes
import hashlib
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
ALL_LEGS = ("infer", "e2e", "cfg4", "cfg5", "longform", "conv", "module", "ref_cuda", "cpu")
TMIX_PARAMS_PER_DIRECTION = 1249792       # SURVEY App. C at C = 512: 5C + 128C + 128C + C + 64C + 64C + C + 4C^2 + 2C
KERNEL_SOURCES = ("wkv6_chunk_lean.cu", "wkv6_chunk_bwd.cu", "wkv6_fwd_tiles.cuh", "wkv6_umma.cuh", "wkv6_common.cuh")


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=50)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", default="b200", choices=["b200", "reference"])
    ap.add_argument("--legs", default=",".join(ALL_LEGS), help="comma-separated subset of: " + ",".join(ALL_LEGS) + " (or 'none')")
    return ap.parse_args()


def peaks():
    try:
        return json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json"))), "measured"
    except Exception:
        return {"hbm_gbs": 6650.0, "bf16_tflops": 1590.0}, "fallback"


def kernel_source_sha():
    h = hashlib.sha256()
    for s in KERNEL_SOURCES:
        with open(os.path.join(ROOT, "paper_accurate_fast_cheap_b200", "csrc", s), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def ncu_traffic():
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the two chunk kernels, from the newest
    profiles/*_ncu_traffic.json (written by scripts/ncu_traffic.py from an `ncu --set full` capture of this bench).
    The file records the hash of the kernel sources it was captured on; a stale capture is not reported."""
    import glob
    files = sorted(glob.glob(os.path.join(ROOT, "profiles", "*_ncu_traffic.json")))
    if not files:
        return {}, "no ncu capture in profiles/"
    rec = json.load(open(files[-1]))
    if rec.get("kernel_source_sha") != kernel_source_sha():
        return {}, f"{os.path.basename(files[-1])} was captured on older kernel sources: not reported"
    return rec.get("traffic", {}), os.path.basename(files[-1])


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
            time.sleep(0.02)

    def summary(self):
        sm = sorted(int(r[0]) for r in self.rows if r and r[0].isdigit())
        mx = [int(r[1]) for r in self.rows if len(r) > 1 and r[1].isdigit()]
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({n for r in self.rows if len(r) >= 6 for n, v in zip(names, r[2:6]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "reasons": reasons, "samples": len(sm)}


def config_dict(args):
    """Identical in both arms (`--impl b200` and `--impl reference`)."""
    c = CFG
    return {"workload": f"configs[1] shape, training step: bi-WKV6 fwd+bwd of a {c['layers']}-block RWKV-6 encoder, "
                        f"d_model={c['C']}, {c['H']} heads x 64, B={c['B']} x T={c['T']} per GPU",
            "layers": c["layers"], "B_per_gpu": c["B"], "T": c["T"], "C": c["C"], "H": c["H"], "mode": "fwd+bwd",
            "parallelism": f"utterance-sharded x{args.gpus}, no data-path collective (gradient all-reduce: see cfg4)",
            "l2": "inputs (1.77 GB per step) exceed the 126 MB L2; every block reads its own tensors"}


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
    os.write(1 if _OUT_FD is None else _OUT_FD, data)


def host_threads():
    return len(os.sched_getaffinity(0)) if hasattr(os, "sched_getaffinity") else (os.cpu_count() or 1)


# ------------------------------------------------------------------------------------------------------------
# CPU port of the reference kernels (oracle/): the reference arm and the cpu_baseline leg
# ------------------------------------------------------------------------------------------------------------
_CPU_SETS = None


def cpu_blocks(nblocks, train=True):
    """`nblocks` bidirectional blocks of the workload (full B x T, fp32) through the oracle port: forward and backward of
    both directions.  Returns seconds."""
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
    for _ in range(nblocks):
        for s, rev in zip(_CPU_SETS, (False, True)):
            oracle.fwd(s["r"], s["k"], s["v"], s["w"], s["u"], rev=rev, prec="f32")
            if train:
                oracle.bwd(s["r"], s["k"], s["v"], s["w"], s["u"], s["gy"], rev=rev, prec="f32")
    return time.perf_counter() - t0


def run_reference(args):
    """Reference arm: the reference has no CPU WKV6 (model.py:105-106 is CUDA-only) and its CUDA kernels cannot be the CPU
    arm, so this times the oracle port of its kernels (fwd + bwd, both directions) on all host cores.  Rank 0 only.
    Every step is a bounded sample: `blocks_per_step` of the 12 blocks (all 12 when the run fits ~3 minutes), and the
    step time is scaled to 12 blocks."""
    if int(os.environ.get("RANK", "0")) != 0:
        return
    import oracle
    oracle.build()
    oracle.set_threads(host_threads())    # torchrun exports OMP_NUM_THREADS=1 to its workers
    c = CFG
    t_block = cpu_blocks(1)
    n_steps = args.steps + args.warmup
    bps = max(1, min(c["layers"], int(180.0 / max(n_steps * t_block, 1e-6))))
    times = []
    for i in range(n_steps):
        dt = cpu_blocks(bps)
        if i >= args.warmup:
            times.append(dt)
    ms = 1e3 * (sum(times) / len(times)) * c["layers"] / bps
    fps = c["B"] * c["T"] / (ms * 1e-3)
    th = oracle.num_threads()
    sample = (f"{len(times)} steps of {bps} of the 12 blocks (both directions, fwd+bwd, B=16 x T=750, fp32), scaled to 12 "
              f"blocks: oracle port of the reference kernels, OpenMP over batch x head on {th} host threads "
              "(the reference itself has no CPU WKV6)")
    line = {"impl": "reference", "metric": "encoder frames/s (bi-WKV6 fwd+bwd)", "value": fps, "unit": "frames/s",
            "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32", "data": "synthetic",
            "config": config_dict(args),
            "cpu_baseline": {"value": fps, "unit": "frames/s", "cores": th, "kind": "port", "sample": sample},
            "e2e": {"value": fps, "unit": "frames/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0}}
    emit(line)


# ------------------------------------------------------------------------------------------------------------
# B200 arm
# ------------------------------------------------------------------------------------------------------------
def make_layer_inputs(torch, dev, layer, dtype, seed_offset=0):
    """Synthetic op-level inputs of SURVEY 8(d) for one block: [r, k, v, w, u, gy] of the L2R set and of the R2L set."""
    c = CFG
    g = torch.Generator(device="cpu").manual_seed(777 + layer + seed_offset)
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
        t.append(torch.randn(B, T, C, generator=g))
        sets.append([x.to(dtype).contiguous().to(dev) for x in t])
    return sets


def bind_to_gpu_numa_node(torch, index):
    """Pinned buffers are first-touched by this process: run it on the cores of the GPU's NUMA node (8 ranks on one
    node's memory was what bent the end-to-end curve in round 1).  Best effort; returns a note for the JSON line."""
    try:
        pr = torch.cuda.get_device_properties(index)
        bdf = f"{pr.pci_domain_id:04x}:{pr.pci_bus_id:02x}:{pr.pci_device_id:02x}.0"
        node = int(open(f"/sys/bus/pci/devices/{bdf}/numa_node").read())
        if node < 0:
            return "gpu numa node unknown"
        cpus = set()
        for part in open(f"/sys/devices/system/node/node{node}/cpulist").read().strip().split(","):
            a, _, b = part.partition("-")
            cpus.update(range(int(a), int(b or a) + 1))
        cpus &= os.sched_getaffinity(0)
        if not cpus:
            return f"node {node}: none of its cores in this process's affinity set"
        os.sched_setaffinity(0, cpus)
        return f"bound to NUMA node {node} ({len(cpus)} cores)"
    except Exception as e:  # noqa: BLE001
        return f"not bound ({type(e).__name__})"


def run_b200(args):
    import torch
    import torch.distributed as dist
    import paper_accurate_fast_cheap_b200 as pk
    from paper_accurate_fast_cheap_b200 import _lib
    from paper_accurate_fast_cheap_b200.sharding import max_over_ranks, shard_rows

    legs = set() if args.legs == "none" else set(x for x in args.legs.split(",") if x)
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    assert torch.cuda.is_available(), "bench.py needs a CUDA device (no CPU fallback)"
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    numa_note = bind_to_gpu_numa_node(torch, local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    L = _lib.lib()
    c = CFG
    B, T, C, H, NL = c["B"], c["T"], c["C"], c["H"], c["layers"]
    dt = torch.bfloat16
    layers = [make_layer_inputs(torch, dev, l, dt, seed_offset=1000 * rank) for l in range(NL)]
    mk = lambda b=B: torch.empty(b, T, C, device=dev, dtype=dt)
    outs = [[mk(), mk()] for _ in range(NL)]
    gouts = [[[mk() for _ in range(4)] + [torch.empty(B, C, device=dev, dtype=dt)] for _ in range(2)] for _ in range(NL)]
    stream = torch.cuda.current_stream().cuda_stream
    BOTH = _lib.BI_L2R | _lib.BI_R2L
    p = lambda t: None if t is None else t.data_ptr()
    ck_bytes = L.wkv6_b200_checkpoint_bytes(B, T, C, H)
    ckpts = [torch.empty((2, max(ck_bytes, 16)), device=dev, dtype=torch.uint8) for _ in range(NL)]
    ws_bytes = L.wkv6_b200_backward_workspace_bytes(0, B, T, C, H, 2)
    ws = torch.empty(max(ws_bytes, 16), device=dev, dtype=torch.uint8)

    def fwd_call(b_, f, b, yf, yb, ckf, ckb, flags=BOTH, st=None):
        """One C-ABI forward launch; a direction whose flag bit is clear gets NULL pointers (it is neither read nor written)."""
        rf, rb = bool(flags & _lib.BI_L2R), bool(flags & _lib.BI_R2L)
        fa = [p(x) for x in f[:5]] if rf else [None] * 5
        ba = [p(x) for x in b[:5]] if rb else [None] * 5
        _lib.check(L.wkv6_b200_bi_forward(0, b_, T, C, H, *fa, None, p(yf) if rf else None, None, p(ckf) if rf else None,
                                          *ba, None, p(yb) if rb else None, None, p(ckb) if rb else None,
                                          flags, None, 0, st if st is not None else stream))

    def bwd_call(b_, f, b, ckf, ckb, gf, gb, flags=BOTH, st=None):
        rf, rb = bool(flags & _lib.BI_L2R), bool(flags & _lib.BI_R2L)
        fa = ([p(x) for x in f[:5]] + [None, p(f[5]), None, p(ckf)] + [p(x) for x in gf] + [None]) if rf else [None] * 15
        ba = ([p(x) for x in b[:5]] + [None, p(b[5]), None, p(ckb)] + [p(x) for x in gb] + [None]) if rb else [None] * 15
        _lib.check(L.wkv6_b200_bi_backward(0, b_, T, C, H, *fa, *ba, flags, p(ws), ws_bytes, st if st is not None else stream))

    def fwd_layer(l, train):
        f, b = layers[l]
        fwd_call(B, f, b, outs[l][0], outs[l][1], ckpts[l][0] if train else None, ckpts[l][1] if train else None)

    def bwd_layer(l):
        f, b = layers[l]
        bwd_call(B, f, b, ckpts[l][0], ckpts[l][1], gouts[l][0], gouts[l][1])

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def allmax(ms):                      # a multi-GPU number is the MAX over ranks of the device-measured time
        return max_over_ranks(ms, dev)

    def region(fn, reps):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        torch.cuda.synchronize()
        e0.record()
        for _ in range(reps):
            fn()
        e1.record()
        torch.cuda.synchronize()
        return e0.elapsed_time(e1) / reps

    def timed(step, steps, warmup, sampler=None, tag=None):
        """`steps` steps, the timed region bracketed by a barrier + synchronize on both sides and by CUDA events on the
        launching stream, nothing but the kernel launches inside (an event record between two launches serialises them
        and defeats the dependent-launch overlap).  Returns ms/step (max over ranks) and the launches in the region."""
        for _ in range(warmup):
            step()
        barrier()
        n0 = pk.launch_count()
        if sampler:
            sampler.start()
        t0, t1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        if tag:
            torch.cuda.nvtx.range_push(tag)       # lets `ncu --nvtx --nvtx-include <tag>/` pick a launch of this very region
        t0.record()
        for _ in range(steps):
            step()
        t1.record()
        if tag:
            torch.cuda.nvtx.range_pop()
        barrier()
        return allmax(t0.elapsed_time(t1)) / steps, pk.launch_count() - n0

    def train_step():
        for l in range(NL):
            fwd_layer(l, True)
        for l in reversed(range(NL)):
            bwd_layer(l)

    def infer_step():
        for l in range(NL):
            fwd_layer(l, False)

    pk_, how = peaks()
    traffic, traffic_src = ncu_traffic()
    fwd_bytes = 2 * 5 * B * T * C * 2                     # both directions: read r,k,v,w + write y, bf16
    bwd_bytes = 2 * 9 * B * T * C * 2                     # read r,k,v,w,gy + write gr,gk,gv,gw

    def roof(kernel, nbytes, ms, traffic_key=None):
        ach = nbytes / (ms * 1e-3) / 1e9
        return {"bound": "hbm", "kernel": kernel, "achieved": ach, "peak": pk_["hbm_gbs"], "unit": "GB/s",
                "frac": ach / pk_["hbm_gbs"], "peak_source": how, "traffic": traffic.get(traffic_key) if traffic_key else None,
                "traffic_source": traffic_src if traffic_key else None, "alg_bytes_per_launch": nbytes, "kernel_ms": ms}

    # ---- headline: the training step (fwd with checkpoint stores + bwd), configs[1] shape
    sampler = ClockSampler(local)
    ms_step, launches = timed(train_step, args.steps, args.warmup, sampler, tag="train_timed")
    t_end = time.time() + 1.5
    while time.time() < t_end and len(sampler.rows) < 6:     # a short timed region: keep the same load up until sampled
        train_step()
        torch.cuda.synchronize()
    sampler.stop_flag = True
    # kernel durations: the step mixes two kernels, so right after the timed region the same launches are timed again as
    # two homogeneous regions (the 12 forward launches, then the 12 backward launches): region time / launches
    ksteps = max(3, min(args.steps, 30))
    fk = region(lambda: [fwd_layer(l, True) for l in range(NL)], ksteps) / NL
    bk = region(lambda: [bwd_layer(l) for l in reversed(range(NL))], ksteps) / NL
    value = world * B * T / (ms_step * 1e-3)
    roofline = roof("wkv6_bwd_chunk_kernel: bidirectional backward, one launch per block (the dominant kernel of the step)",
                    bwd_bytes, bk, "bwd")
    extra = {"roofline_fwd": roof("wkv6_fwd_chunk_lean_kernel (+ checkpoint stores): bidirectional forward, one launch per block",
                                  fwd_bytes, fk, "fwd_train"),
             "roofline_fwd_bwd": roof("forward + backward launches of one block", fwd_bytes + bwd_bytes, fk + bk),
             "pinned_memory": numa_note}

    def leg(name, fn):
        if name not in legs:
            return
        try:
            r = fn()
            if r is not None:
                extra[name] = r
        except Exception as e:  # noqa: BLE001  (a failing side leg must not take the headline line down)
            extra[name] = {"error": f"{type(e).__name__}: {e}"[:300]}
            torch.cuda.synchronize()

    # ---- inference (configs[1] literally): the forward alone, no checkpoint stores
    def leg_infer():
        ms, n = timed(infer_step, args.steps, args.warmup, tag="infer_timed")
        return {"metric": "encoder frames/s (bi-WKV6 fwd)", "value": world * B * T / (ms * 1e-3), "unit": "frames/s",
                "ms_per_step": ms, "steps": args.steps, "gpu_launches": n,
                "roofline": roof("wkv6_fwd_chunk_lean_kernel: bidirectional forward, one launch per block", fwd_bytes, ms / NL, "fwd")}

    # ---- configs[3]: B = 8 per GPU, direction dropout, NCCL all-reduce of the Tmix gradients
    def leg_cfg4():
        B4 = 8
        # seeded CPU Bernoulli sequence (rwkv_wrapper_bidirectional_direction_dropout_both.py:53-71), same on every rank
        from torch.distributions.bernoulli import Bernoulli
        cpu_state = torch.get_rng_state()
        torch.manual_seed(777)
        n4 = max(3, min(args.steps, 30))
        seq = []
        for _ in range((3 + n4) * NL):
            if Bernoulli(0.8).sample((1,)) == 1:
                seq.append(BOTH)
            else:
                seq.append(_lib.BI_L2R if Bernoulli(0.8).sample((1,)) <= 0.5 else _lib.BI_R2L)
        torch.set_rng_state(cpu_state)
        NP = TMIX_PARAMS_PER_DIRECTION
        # buckets in the order the backward fills them (block 11 first), DDP-sized (bucket_cap_mb = 25: up to 5 blocks = 25 MB),
        # shrinking towards the end: the bucket that closes last cannot overlap any backward work, so it is the smallest
        BUCKET_BLOCKS = [5, 4, 2, 1] if NL == 12 else [NL]
        NBK = len(BUCKET_BLOCKS)
        first_block, bucket_of, pos_in = [], {}, {}
        hi = NL
        for k_, nb_ in enumerate(BUCKET_BLOCKS):
            lo = hi - nb_
            first_block.append(lo)
            for l in range(lo, hi):
                bucket_of[l], pos_in[l] = k_, l - lo
            hi = lo
        gb = torch.Generator(device=dev).manual_seed(4242 + rank)
        # the surrounding model's gradients (GEMM / LoRA / LayerNorm weights) are outside the path: synthetic, fixed
        synth = [torch.randn(2 * NP, device=dev, generator=gb).mul_(1e-2).to(dt) for _ in range(NL)]
        buckets = [torch.cat([synth[l] for l in range(first_block[k], first_block[k] + BUCKET_BLOCKS[k])]) for k in range(NBK)]
        half_of = lambda l, d: buckets[bucket_of[l]][(pos_in[l] * 2 + d) * NP:(pos_in[l] * 2 + d + 1) * NP]
        zeroed = [[False, False] for _ in range(NL)]
        U_OFF = 5 * C + 2 * 128 * C + C + 2 * 64 * C                     # time_faaaa inside one direction's flat gradient
        comm = torch.cuda.Stream()
        evs = [torch.cuda.Event() for _ in range(NBK)]
        cursor = [0]

        def step(collective):
            i0 = cursor[0]
            cursor[0] = (cursor[0] + NL) % len(seq)
            fl = seq[i0:i0 + NL]
            for l in range(NL):
                f, b = layers[l]
                fwd_call(B4, [x[:B4] if x.dim() == 3 else x for x in f], [x[:B4] if x.dim() == 3 else x for x in b],
                         outs[l][0][:B4], outs[l][1][:B4], ckpts[l][0], ckpts[l][1], fl[l])
            works = []
            for l in reversed(range(NL)):
                f, b = layers[l]
                gf = [g[:B4] for g in gouts[l][0]]
                gb_ = [g[:B4] for g in gouts[l][1]]
                bwd_call(B4, [x[:B4] if x.dim() == 3 else x for x in f], [x[:B4] if x.dim() == 3 else x for x in b],
                         ckpts[l][0], ckpts[l][1], gf, gb_, fl[l])
                # the block's slice of its gradient bucket: this launch's gu.sum(0) (model.py:151) goes into time_faaaa's
                # slot; a dropped direction's half is zero, which is what DDP(find_unused_parameters=True) reduces for
                # parameters that got no gradient (the all-reduce volume does not depend on the dropout outcome)
                for d, (g5, bit) in enumerate(((gf[4], _lib.BI_L2R), (gb_[4], _lib.BI_R2L))):
                    half = half_of(l, d)
                    if fl[l] & bit:
                        if zeroed[l][d]:
                            half.copy_(synth[l][d * NP:(d + 1) * NP])
                            zeroed[l][d] = False
                        torch.sum(g5, 0, out=half[U_OFF:U_OFF + C])
                    elif not zeroed[l][d]:
                        half.zero_()
                        zeroed[l][d] = True
                if collective and world > 1 and pos_in[l] == 0:        # the bucket's last block (backward order) is done
                    k_ = bucket_of[l]
                    evs[k_].record()
                    with torch.cuda.stream(comm):
                        comm.wait_event(evs[k_])
                        works.append(dist.all_reduce(buckets[k_], op=dist.ReduceOp.AVG, async_op=True))
            for w_ in works:
                w_.wait()
            if works:
                torch.cuda.current_stream().wait_stream(comm)

        def run(collective):
            cursor[0] = 0
            return timed(lambda: step(collective), n4, 3)

        ms_with, n_l = run(True)
        ms_without, _ = run(False)
        res = {"metric": "encoder frames/s (bi-WKV6 fwd+bwd, direction dropout, gradient all-reduce)",
               "workload": f"configs[3]: {B4} utterances x T={T} per GPU x {world} GPUs, 12 blocks, bf16, direction dropout p=0.2 "
                           f"(dir_layer_drop_both), bf16 gradient buckets of {BUCKET_BLOCKS} blocks x 2 x 1 249 792",
               "value": world * B4 * T / (ms_with * 1e-3), "unit": "frames/s", "ms_per_step": ms_with,
               "ms_per_step_without_collective": ms_without, "steps": n4, "gpu_launches": n_l,
               "directions_run": {"both": seq[:n4 * NL].count(BOTH), "l2r_only": seq[:n4 * NL].count(_lib.BI_L2R),
                                  "r2l_only": seq[:n4 * NL].count(_lib.BI_R2L)},
               "allreduce_bytes_per_step": NL * 2 * NP * 2}
        if world > 1:
            def only_comm():
                ws_ = [dist.all_reduce(bk_, op=dist.ReduceOp.AVG, async_op=True) for bk_ in buckets]
                for w_ in ws_:
                    w_.wait()
            for _ in range(2):
                only_comm()
            barrier()
            alone = allmax(region(only_comm, 5))
            exposed = max(ms_with - ms_without, 0.0)
            # post-reduce equality across ranks: every rank must hold bit-identical averaged buckets
            step(True)
            torch.cuda.synchronize()
            sig = torch.stack([bk_.view(torch.int16).to(torch.int64).sum() for bk_ in buckets])
            allsig = [torch.empty_like(sig) for _ in range(world)]
            dist.all_gather(allsig, sig)
            same = all(bool((s_ == allsig[0]).all()) for s_ in allsig)
            assert same, "gradient buckets differ across ranks after the all-reduce"
            res.update({"collective": f"NCCL all_reduce(AVG) bf16, {NBK} buckets of {BUCKET_BLOCKS} blocks on a side stream, each issued "
                                      "right after the backward launch of its last block", "allreduce_alone_ms": alone, "exposed_ms": exposed,
                        "overlap_fraction": max(0.0, min(1.0, 1.0 - exposed / max(alone, 1e-9))),
                        "buckets_identical_across_ranks": same})
        else:
            res["collective"] = "none (1 rank)"
        return res

    # ---- configs[4]: fp32, one 1-hour utterance per GPU, d_model 1024, 16 heads, one bidirectional block, forward
    def leg_cfg5():
        T5, C5, H5 = 90000, 1024, 16
        g5 = torch.Generator(device=dev).manual_seed(555 + rank)
        sets = []
        for _ in range(2):
            t = [torch.randn(1, T5, C5, device=dev, generator=g5) for _ in range(3)]
            t.append(-6 + 5 * torch.rand(1, T5, C5, device=dev, generator=g5))
            t.append(0.3 * torch.randn(H5, 64, device=dev, generator=g5))
            sets.append(t)
        ys = [torch.empty(1, T5, C5, device=dev) for _ in range(2)]
        nb = L.wkv6_b200_forward_scratch_bytes(1, 1, T5, C5, H5, 2)
        scratch = torch.empty(max(nb, 16), device=dev, dtype=torch.uint8)
        f, b = sets

        def once():
            _lib.check(L.wkv6_b200_bi_forward(1, 1, T5, C5, H5, *[p(x) for x in f], None, p(ys[0]), None, None,
                                              *[p(x) for x in b], None, p(ys[1]), None, None, BOTH,
                                              p(scratch) if nb else None, nb, stream))
        n0 = pk.launch_count()
        once()
        per = pk.launch_count() - n0
        ms, _ = timed(once, 5, 2)
        nbytes = 2 * 5 * T5 * C5 * 4
        flops = 2 * 7 * 64 * C5 * T5                                       # 7 N^2 per (head, step), both directions
        ffma_peak = 148 * 128 * 2 * 1.965e9
        hbm = nbytes / (ms * 1e-3) / 1e9
        return {"metric": "encoder frames/s (bi-WKV6 fwd, fp32)", "value": world * T5 / (ms * 1e-3), "unit": "frames/s",
                "workload": f"configs[4]: fp32, 1 utterance x T={T5} per GPU, d_model={C5}, {H5} heads, one bidirectional block, fwd",
                "ms_per_block": ms, "launches_per_block": per,
                "roofline": {"bound": "fp32 FMA pipe (exact recurrent form: the 1e-5 tolerance rules out bf16 / tf32 operands)",
                             "achieved": flops / (ms * 1e-3) / 1e12, "peak": ffma_peak / 1e12, "unit": "TFLOP/s",
                             "frac": flops / (ms * 1e-3) / ffma_peak,
                             "peak_source": "148 SMs x 128 lanes x 2 x 1.965 GHz (derived, BASELINE.md section 2)",
                             "algorithmic_flops": flops,
                             "note": "the time-segmented launch runs the recurrence twice (local states, then outputs): "
                                     "2x these flops are issued"},
                "hbm": {"achieved": hbm, "peak": pk_["hbm_gbs"], "unit": "GB/s", "frac": hbm / pk_["hbm_gbs"],
                        "alg_bytes_per_launch": nbytes}}

    # ---- configs[2]: 8 utterances of T = 30 000 in total, sharded by utterance (8 / N per GPU: strong scaling), forward
    def leg_longform():
        if world > 8:
            return None
        _lb, _le = shard_rows(8, rank, world)          # configs[2]: 8 utterances sharded over the ranks
        LT, LB = 30000, max(1, _le - _lb)
        gl = torch.Generator(device=dev).manual_seed(777 + rank)
        lsets = []
        for _ in range(2):
            t = [torch.randn(LB, LT, C, device=dev, generator=gl).to(dt) for _ in range(3)]
            t.append((-6 + 5 * torch.rand(LB, LT, C, device=dev, generator=gl)).to(dt))
            t.append((0.3 * torch.randn(H, 64, device=dev, generator=gl)).to(dt))
            lsets.append(t)
        ly = [torch.empty(LB, LT, C, device=dev, dtype=dt) for _ in range(2)]
        nb = L.wkv6_b200_forward_scratch_bytes(0, LB, LT, C, H, 2)
        scratch = torch.empty(max(nb, 16), device=dev, dtype=torch.uint8)
        f, b = lsets

        def lstep():
            for _ in range(NL):
                _lib.check(L.wkv6_b200_bi_forward(0, LB, LT, C, H, *[p(x) for x in f], None, p(ly[0]), None, None,
                                                  *[p(x) for x in b], None, p(ly[1]), None, None, BOTH,
                                                  p(scratch) if nb else None, nb, stream))
        lms, n = timed(lstep, 3, 2)
        nbytes = 2 * 5 * LB * LT * C * 2
        # long-form fine-tuning shape (conf/rwkv/*FT-LFXL*.yaml: 50 000 input frames -> T' = 12 499), one utterance per GPU,
        # fwd (checkpoints) + bwd of one block: both time-segmented (the reference's backward cannot run T > ctx_len at all)
        FT = 12499
        fsets = [[x[:1, :FT].contiguous() if x.dim() == 3 else x for x in s_] for s_ in lsets]
        fgy = [torch.randn(1, FT, C, device=dev, generator=gl).to(dt) for _ in range(2)]
        fy = [torch.empty(1, FT, C, device=dev, dtype=dt) for _ in range(2)]
        fg = [[torch.empty(1, FT, C, device=dev, dtype=dt) for _ in range(4)] + [torch.empty(1, C, device=dev, dtype=dt)] for _ in range(2)]
        fck = torch.empty((2, L.wkv6_b200_checkpoint_bytes(1, FT, C, H)), device=dev, dtype=torch.uint8)
        fnb = L.wkv6_b200_forward_scratch_bytes(0, 1, FT, C, H, 2)
        bnb = L.wkv6_b200_backward_scratch_bytes(0, 1, FT, C, H, 2)
        fsc = torch.empty(max(fnb, bnb, 16), device=dev, dtype=torch.uint8)

        def ft_step(seg):
            ff, fb = fsets
            _lib.check(L.wkv6_b200_bi_forward(0, 1, FT, C, H, *[p(x) for x in ff], None, p(fy[0]), None, p(fck[0]),
                                              *[p(x) for x in fb], None, p(fy[1]), None, p(fck[1]), BOTH,
                                              p(fsc) if seg else None, fnb if seg else 0, stream))
            _lib.check(L.wkv6_b200_bi_backward(0, 1, FT, C, H,
                                               *[p(x) for x in ff], None, p(fgy[0]), None, p(fck[0]), *[p(x) for x in fg[0]], None,
                                               *[p(x) for x in fb], None, p(fgy[1]), None, p(fck[1]), *[p(x) for x in fg[1]], None,
                                               BOTH, p(fsc) if seg else None, bnb if seg else 0, stream))
        ft_seg, ft_n = timed(lambda: ft_step(True), 5, 2)
        ft_plain, _ = timed(lambda: ft_step(False), 3, 1)
        ft = {"workload": f"1 utterance x T={FT} per GPU, C=512, 8 heads, bf16, ONE block, fwd (checkpoints) + bwd, both directions",
              "ms_per_block": ft_seg, "ms_per_block_one_cta_per_sequence": ft_plain, "launches_per_block": ft_n // 5,
              "frames_per_s_per_block": world * FT / (ft_seg * 1e-3)}
        return {"finetune_longform": ft,"metric": "encoder frames/s (bi-WKV6 fwd)", "value": world * LB * LT / (lms * 1e-3), "unit": "frames/s",
                "workload": f"configs[2]: 8 utterances x T=30000 in total, {LB} per GPU, C=512, 8 heads, bf16, 12 blocks, fwd",
                "scaling": "strong", "ms_per_step": lms, "steps": 3, "launches_per_block": n // (3 * NL),
                "hbm_frac": nbytes / (lms / NL * 1e-3) / 1e9 / pk_["hbm_gbs"]}

    # ---- the reference's own CUDA kernels on this GPU (wkv6_cuda.cu unmodified; one direction per launch, as the reference
    # runs them, plus the two flips + add + cast its wrapper adds per block: rwkv_wrapper_bidirectional.py:43-56)
    def leg_ref_cuda():
        so = os.path.join(ROOT, "oracle", "_ref", "libwkv6_ref.so")
        if rank != 0 or not os.path.exists(so):
            return None if rank != 0 else {"unavailable": "oracle/_ref/libwkv6_ref.so not built"}
        ref = ctypes.CDLL(so)
        f = layers[0][0]
        y = outs[0][0]
        g = gouts[0][0]
        vp = lambda t: ctypes.c_void_p(t.data_ptr())

        def fwd():
            rc = ref.wkv6_ref_forward(0, B, T, C, H, vp(f[0]), vp(f[1]), vp(f[2]), vp(f[3]), vp(f[4]), vp(y))
            assert rc == 0, f"reference forward: cuda error {rc}"

        def bwd():
            rc = ref.wkv6_ref_backward(0, B, T, C, H, vp(f[0]), vp(f[1]), vp(f[2]), vp(f[3]), vp(f[4]), vp(f[5]), *[vp(x) for x in g])
            assert rc == 0, f"reference backward: error {rc}"

        def wrapper():
            xf = torch.flip(f[0], [1])
            yb = torch.flip(y, [1])
            return ((y + yb) / 2).float(), xf

        with torch.cuda.stream(torch.cuda.default_stream()):       # the reference launches on the legacy default stream
            for fn in (fwd, bwd, wrapper):
                fn(); fn()
            t_f, t_b, t_w = region(fwd, 5), region(bwd, 3), region(wrapper, 5)
        step_ms = NL * (2 * (t_f + t_b) + 2 * t_w)
        return {"what": "reference CUDA kernels (wenet/rwkv_v6/cuda/wkv6_cuda.cu, unmodified, the reference's JIT flags, sm_100a) "
                        "on this GPU: one direction per launch", "fwd_ms_per_direction": t_f, "bwd_ms_per_direction": t_b,
                "wrapper_flip_add_cast_ms": t_w, "fwd_hbm_frac": (fwd_bytes / 2) / (t_f * 1e-3) / 1e9 / pk_["hbm_gbs"],
                "bwd_hbm_frac": (bwd_bytes / 2) / (t_b * 1e-3) / 1e9 / pk_["hbm_gbs"],
                "metric": "encoder frames/s (bi-WKV6 fwd+bwd)", "value": B * T / (step_ms * 1e-3), "unit": "frames/s",
                "ms_per_step": step_ms, "speedup_of_this_repo": step_ms / ms_step,
                "kernel_speedup_fwd": 2 * t_f / fk, "kernel_speedup_bwd": 2 * t_b / bk}

    # ---- end to end: the headline step through the C ABI with HOST buffers.  Per block: one pinned slab (r,k,v,w,u,gy of
    # both directions) -> one H2D copy -> fwd -> bwd -> one D2H copy of the slab of gradients (gr,gk,gv,gw,gu of both)
    def leg_e2e():
        al = lambda n: (n + 127) // 128 * 128                 # element counts rounded to 256 bytes
        n_big, n_u, n_gu = B * T * C, al(H * 64), al(B * C)
        in_elems = 2 * (5 * n_big + n_u)
        out_elems = 2 * (4 * n_big + n_gu)

        def carve_in(slab):
            sets, o = [], 0
            for _ in range(2):
                t = []
                for i in range(6):
                    n = n_u if i == 4 else n_big
                    t.append(slab[o:o + (H * 64 if i == 4 else n_big)].view((H, 64) if i == 4 else (B, T, C)))
                    o += n
                sets.append(t)
            return sets

        def carve_out(slab):
            sets, o = [], 0
            for _ in range(2):
                t = []
                for i in range(5):
                    t.append(slab[o:o + (B * C if i == 4 else n_big)].view((B, C) if i == 4 else (B, T, C)))
                    o += n_gu if i == 4 else n_big
                sets.append(t)
            return sets

        hin = [torch.empty(in_elems, dtype=dt).pin_memory() for _ in range(2)]          # 2 distinct blocks, reused
        for k_, h_ in enumerate(hin):
            for src, dst in zip(layers[k_], carve_in(h_)):
                for a_, b__ in zip(src, dst):
                    b__.copy_(a_)
        hout = [torch.empty(out_elems, dtype=dt).pin_memory() for _ in range(2)]
        din = [torch.empty(in_elems, device=dev, dtype=dt) for _ in range(2)]
        dout = [torch.empty(out_elems, device=dev, dtype=dt) for _ in range(2)]
        din_v, dout_v = [carve_in(s) for s in din], [carve_out(s) for s in dout]
        s_in, s_out = torch.cuda.Stream(), torch.cuda.Stream()
        s_cmp = torch.cuda.current_stream()
        ev_in = [torch.cuda.Event() for _ in range(2)]
        ev_cmp = [torch.cuda.Event() for _ in range(2)]
        ev_out = [torch.cuda.Event() for _ in range(2)]

        def e2e_step():
            # three streams, double-buffered staging: H2D of block l+1 and D2H of block l-1 overlap the kernels of block l
            for l in range(NL):
                k_ = l & 1
                with torch.cuda.stream(s_in):
                    s_in.wait_event(ev_cmp[k_])              # the kernels that last read this staging slot are done
                    din[k_].copy_(hin[k_], non_blocking=True)
                    ev_in[k_].record(s_in)
                s_cmp.wait_event(ev_in[k_])
                s_cmp.wait_event(ev_out[k_])                 # gradient staging slot k has been copied back
                f, b = din_v[k_]
                fwd_call(B, f, b, outs[l][0], outs[l][1], ckpts[l][0], ckpts[l][1])
                bwd_call(B, f, b, ckpts[l][0], ckpts[l][1], dout_v[k_][0], dout_v[k_][1])
                ev_cmp[k_].record(s_cmp)
                with torch.cuda.stream(s_out):
                    s_out.wait_event(ev_cmp[k_])
                    hout[k_].copy_(dout[k_], non_blocking=True)
                    ev_out[k_].record(s_out)
            s_cmp.wait_stream(s_out)                         # the step ends when its last gradients are on the host

        n_e = max(2, min(args.steps, 5))
        ms, _ = timed(e2e_step, n_e, 1)
        return {"value": world * B * T / (ms * 1e-3), "unit": "frames/s", "h2d_bytes_per_step": in_elems * 2 * NL,
                "d2h_bytes_per_step": out_elems * 2 * NL, "steps": n_e, "ms_per_step": ms,
                "what": "fwd+bwd through the C ABI from pinned host slabs: H2D of r,k,v,w,u,gy and D2H of gr,gk,gv,gw,gu of both "
                        "directions inside the timed region, every block"}

    # ---- ConvolutionModule (wenet/transformer/convolution.py:89-144), kernel 31 / layer_norm / swish, fp32
    def leg_conv():
        from paper_accurate_fast_cheap_b200.convolution import ConvolutionModule, convmod_core
        K = 31
        torch.manual_seed(5)
        mod = ConvolutionModule(C, K, torch.nn.SiLU(), "layer_norm", False, True).to(dev).eval()
        gz = torch.Generator(device=dev).manual_seed(99 + rank)
        zs = [torch.randn(B, T, 2 * C, device=dev, generator=gz) for _ in range(4)]      # 4 x 49 MB: every launch reads cold z
        xs = [torch.randn(B, T, C, device=dev, generator=gz) for _ in range(4)]
        dw, nm = mod.depthwise_conv, mod.norm
        cur = [0]

        def core():
            cur[0] = (cur[0] + 1) & 3
            return convmod_core(zs[cur[0]], dw.weight, dw.bias, nm.weight, nm.bias, nm.eps, T, (K - 1) // 2)

        def core_eager():          # the reference's ops between the two pointwise convolutions (:121-137), on [B, 2C, T]
            cur[0] = (cur[0] + 1) & 3
            z = zs[cur[0]].transpose(1, 2)
            y = dw(torch.nn.functional.glu(z, dim=1)).transpose(1, 2)
            return torch.nn.functional.silu(nm(y)).transpose(1, 2)

        def module(fused):
            cur[0] = (cur[0] + 1) & 3
            with torch.no_grad():
                return mod(xs[cur[0]]) if fused else mod._forward_eager(xs[cur[0]], torch.ones((0, 0, 0), dtype=torch.bool, device=dev),
                                                                          torch.zeros((0, 0, 0), device=dev))
        with torch.no_grad():
            for fn in (core, core_eager):
                fn(); fn()
            module(True); module(False)
            reps = max(5, min(args.steps, 50))
            n0 = pk.launch_count()
            t_core = allmax(region(core, reps))
            n_l = pk.launch_count() - n0
            t_eager = allmax(region(core_eager, reps))
            t_mod, t_mod_eager = allmax(region(lambda: module(True), reps)), allmax(region(lambda: module(False), reps))
        nbytes = B * T * 3 * C * 4
        ach = nbytes / (t_core * 1e-3) / 1e9
        return {"metric": "encoder frames/s (ConvolutionModule core: GLU + depthwise k=31 + LayerNorm + swish)",
                "workload": f"configs[1] shape, one block: B={B} x T={T}, d_model={C}, kernel 31, fp32", "unit": "frames/s",
                "value": world * B * T / (t_core * 1e-3), "ms_per_block": t_core, "gpu_launches": n_l,
                "eager_ops_ms_per_block": t_eager, "speedup_vs_eager_ops": t_eager / t_core,
                "module_ms_per_block": t_mod, "module_eager_ms_per_block": t_mod_eager,
                "roofline": {"bound": "hbm", "kernel": "convmod_core_kernel<float, 31>", "achieved": ach, "peak": pk_["hbm_gbs"],
                             "unit": "GB/s", "frac": ach / pk_["hbm_gbs"], "peak_source": how,
                             "alg_bytes_per_launch": nbytes, "kernel_ms": t_core,
                             "note": "algorithmic bytes: read z [B,T,2C], write y [B,T,C]"}}

    # ---- module level: what wenet's encoder layer calls (self_attn(x, x, x, mask, pos_emb, cache)), 12 blocks
    def leg_module():
        import paper_accurate_fast_cheap_b200.wrappers as W
        so = os.path.join(ROOT, "oracle", "_ref", "libwkv6_ref.so")
        ref = ctypes.CDLL(so) if os.path.exists(so) else None
        vp = lambda t: ctypes.c_void_p(t.data_ptr())

        class RefWKV(torch.autograd.Function):           # model.py:108-159 on the reference's own kernels
            @staticmethod
            def forward(ctx, r, k, v, w, u):
                Bq, Tq, Cq = r.shape
                ctx.save_for_backward(r, k, v, w, u)
                y = torch.empty_like(r)
                assert ref.wkv6_ref_forward(0, Bq, Tq, Cq, Cq // 64, vp(r), vp(k), vp(v), vp(w), vp(u), vp(y)) == 0
                return y

            @staticmethod
            def backward(ctx, gy):
                r, k, v, w, u = ctx.saved_tensors
                Bq, Tq, Cq = r.shape
                gy = gy.contiguous()
                g4 = [torch.empty_like(r) for _ in range(4)]
                gu = torch.empty((Bq, Cq), device=r.device, dtype=r.dtype)
                assert ref.wkv6_ref_backward(0, Bq, Tq, Cq, Cq // 64, vp(r), vp(k), vp(v), vp(w), vp(u), vp(gy), *[vp(t) for t in g4], vp(gu)) == 0
                return (*g4, torch.sum(gu, 0).view(Cq // 64, 64))

        def ref_tmix(blk, x):                            # RWKV_Tmix_x060c.forward as the reference writes it (model.py:270-325)
            Bq, Tq, Cq = x.size()
            xx = torch.nn.functional.pad(x[:, :-1], (0, 0, 1, 0)) - x
            xxx = x + xx * blk.time_maa_x
            xxx = torch.tanh(xxx @ blk.time_maa_rkvw_w1).view(Bq * Tq, 4, -1).transpose(0, 1)
            xxx = torch.bmm(xxx, blk.time_maa_rkvw_w2).view(4, Bq, Tq, Cq)
            mr, mk, mv, mw = xxx.unbind(dim=0)
            r = blk.receptance(x + xx * (blk.time_maa_r + mr))
            k = blk.key(x + xx * (blk.time_maa_k + mk))
            v = blk.value(x + xx * (blk.time_maa_v + mv))
            ww = x + xx * (blk.time_maa_w + mw)
            w = blk.time_decay + torch.tanh(ww @ blk.time_decay_w1) @ blk.time_decay_w2
            y = RefWKV.apply(r.contiguous(), k.contiguous(), v.contiguous(), w.contiguous(), blk.time_faaaa)
            return blk.output(blk.ln_x(y))

        def ref_bi(mod, x, run_f=True, run_b=True):      # rwkv_wrapper_bidirectional.py:43-56 / ..._dropout_both.py:53-99
            xb = x.to(torch.bfloat16)
            f, b = mod._blocks()
            o_f = ref_tmix(f, xb) if run_f else None
            o_b = torch.flip(ref_tmix(b, torch.flip(xb, [1])), [1]) if run_b else None
            if run_f and run_b:
                return ((o_f + o_b) / 2).float()
            return (o_f if run_f else o_b).float()

        torch.manual_seed(777)
        infer_mods = [W.RWKV_TmixWrapper_bidirectional(64, C, NL, "v6", "bi", 4096, True, l).to(dev).eval() for l in range(NL)]
        train_mods = [W.RWKV_TmixWrapper_bidirectional_direction_dropout_both(64, C, NL, "v6", "bi", 4096, True, l).to(dev).train()
                      for l in range(NL)]
        with torch.no_grad():                            # fresh modules have zero LoRA-A matrices: make every GEMM do real work
            for m_ in infer_mods + train_mods:
                for blk in m_._blocks():
                    blk.time_maa_rkvw_w1.normal_(0, 0.02)
                    blk.time_decay_w1.normal_(0, 0.02)
        xg = torch.Generator(device=dev).manual_seed(31 + rank)
        x16 = torch.randn(B, T, C, device=dev, generator=xg)
        x8 = x16[:8].contiguous()
        res = {"workload": f"12 attention-slot modules, d_model={C}, bf16 inside / fp32 interface, T={T}: inference B={B} "
                           "(RWKV_TmixWrapper_bidirectional), training step B=8 (..._direction_dropout_both, p=0.2, same seeded draws)"}

        # every block sees the same input (no residual stream / feed-forward between them here: chained, the 12 bare blocks
        # amplify a bf16-rounding difference to O(1) and outputs could not be compared)
        def infer_ours():
            with torch.no_grad():
                return torch.stack([m_(x16, x16, x16)[0] for m_ in infer_mods])

        def infer_ref():
            with torch.no_grad():
                return torch.stack([ref_bi(m_, x16) for m_ in infer_mods])

        def train_step(ours):
            torch.manual_seed(4321)                      # both variants see the same Bernoulli sequence
            h = x8.clone().requires_grad_(True)
            loss = 0.0
            for m_ in train_mods:
                if ours:
                    o = m_(h, h, h)[0]
                else:
                    run_f, run_b = m_._directions()
                    o = ref_bi(m_, h, run_f, run_b)
                loss = loss + o.float().square().mean()
            loss.backward()
            for m_ in train_mods:
                m_.zero_grad(set_to_none=True)
            return h.grad

        reps = max(3, min(args.steps, 10))
        a_ = infer_ours()
        t_io = allmax(region(infer_ours, reps))
        res["infer"] = {"ms_per_step": t_io, "frames_per_s": world * B * T / (t_io * 1e-3)}
        graphed = W.GraphedInference(infer_mods, x16)          # the same 12 modules as one CUDA graph replay
        g_out = torch.stack(graphed(x16))
        t_ig = allmax(region(lambda: graphed(x16), reps))
        res["infer"].update({"cuda_graph_ms_per_step": t_ig, "cuda_graph_frames_per_s": world * B * T / (t_ig * 1e-3),
                             "cuda_graph_equals_eager": bool(torch.equal(g_out, a_))})
        g_o = train_step(True)
        t_to = allmax(region(lambda: train_step(True), reps))
        res["train"] = {"ms_per_step": t_to, "frames_per_s": world * 8 * T / (t_to * 1e-3)}
        slots = [W.GraphedTraining(m_, x8) for m_ in train_mods]     # forward + backward graphs per direction outcome

        def train_step_graphed():
            torch.manual_seed(4321)
            h = x8.clone().requires_grad_(True)
            loss = 0.0
            for sl in slots:
                loss = loss + sl(h).float().square().mean()
            loss.backward()
            for m_ in train_mods:
                m_.zero_grad(set_to_none=True)
            return h.grad

        for _ in range(4):                                # captures the outcomes of the seeded sequence
            g_g = train_step_graphed()
        t_tg = allmax(region(train_step_graphed, reps))
        res["train"].update({"cuda_graph_ms_per_step": t_tg, "cuda_graph_frames_per_s": world * 8 * T / (t_tg * 1e-3),
                             "cuda_graph_equals_eager": bool(torch.equal(g_g, g_o))})
        if ref is not None and rank == 0:
            with torch.cuda.stream(torch.cuda.default_stream()):     # the reference's launchers use the legacy default stream
                b_ = infer_ref()
                t_ir = region(infer_ref, reps)
                g_r = train_step(False)
                t_tr = region(lambda: train_step(False), max(2, reps // 2))
            rel = lambda p_, q_: float((p_.double() - q_.double()).abs().max() / q_.double().abs().max())
            res["infer"].update({"reference_formulation_ms": t_ir, "speedup": t_ir / t_io, "speedup_cuda_graph": t_ir / t_ig,
                                 "max_rel_diff_of_outputs": rel(a_, b_)})
            res["train"].update({"reference_formulation_ms": t_tr, "speedup": t_tr / t_to, "speedup_cuda_graph": t_tr / t_tg,
                                 "max_rel_diff_of_input_grads": rel(g_o, g_r)})
        return res

    leg("infer", leg_infer)
    leg("conv", leg_conv)
    leg("module", leg_module)
    leg("cfg4", leg_cfg4)
    leg("longform", leg_longform)
    leg("cfg5", leg_cfg5)
    leg("ref_cuda", leg_ref_cuda)
    leg("e2e", leg_e2e)
    e2e = extra.pop("e2e", None)

    if rank == 0:
        cpu = None
        if world == 1 and "cpu" in legs:
            import oracle
            oracle.set_threads(host_threads())
            t1 = cpu_blocks(1)
            n = max(1, min(100, int(12.0 / max(t1, 1e-3))))
            secs = cpu_blocks(n)
            fps = B * T / (secs / n * NL)
            cpu = {"value": fps, "unit": "frames/s", "cores": oracle.num_threads(), "kind": "port",
                   "sample": f"{n} of the workload's blocks (both directions, fwd+bwd, B=16 x T=750, fp32), scaled to 12 blocks: "
                             f"{secs:.1f} s of CPU work on {oracle.num_threads()} threads"}
        line = {"metric": "encoder frames/s (bi-WKV6 fwd+bwd)", "value": value, "unit": "frames/s",
                "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms_step, "higher_is_better": True,
                "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic",
                "config": config_dict(args), "roofline": roofline, "cpu_baseline": cpu, "e2e": e2e,
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
