"""Turns `ncu -i x.ncu-rep --page raw --csv` dumps of the chunk kernels (captured from bench.py, see scripts/gpu_prof_r2.sh)
into profiles/<tag>_ncu_traffic.json: dram__bytes_read.sum + dram__bytes_write.sum per launch, keyed the way bench.py's
`roofline.traffic` looks them up, together with the hash of the kernel sources the capture was taken on (bench.py does not
report a capture that is older than the sources) and a summary text of the counters DESIGN.md quotes.

    python scripts/ncu_traffic.py <tag> fwd=<raw.csv> fwd_train=<raw.csv> bwd=<raw.csv>
"""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed",
        "smsp__average_warps_issue_stalled_no_instruction_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__inst_executed.sum", "launch__registers_per_thread", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "launch__waves_per_multiprocessor", "launch__grid_size"]


def to_bytes(v, unit):
    v = float(v.replace(",", ""))
    return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(unit, 1)


def read_raw(path):
    rows = list(csv.reader(open(path)))
    hdr = next(i for i, r in enumerate(rows) if r and r[0] == "ID")
    names, units, vals = rows[hdr], rows[hdr + 1], rows[hdr + 2]
    return {n: (v, u) for n, u, v in zip(names, units, vals)}


def main():
    tag = sys.argv[1]
    traffic, text = {}, []
    for arg in sys.argv[2:]:
        key, path = arg.split("=", 1)
        m = read_raw(path)
        rd, wr = to_bytes(*m["dram__bytes_read.sum"]), to_bytes(*m["dram__bytes_write.sum"])
        traffic[key] = rd + wr
        text.append(f"== {key}: {m['Kernel Name'][0]}  (grid {m.get('launch__grid_size', ('?', ''))[0]})")
        for n in WANT:
            if n in m:
                text.append(f"  {n} [{m[n][1]}] = {m[n][0]}")
    head = subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True, cwd=ROOT).stdout.strip()
    out = {"traffic": traffic, "unit": "bytes per launch (dram__bytes_read.sum + dram__bytes_write.sum)",
           "kernel_source_sha": bench.kernel_source_sha(), "captured_at_commit": head,
           "command": "ncu --set full --clock-control none on python bench.py --legs infer --steps 2 --warmup 1 (scripts/gpu_prof_r2.sh)"}
    json.dump(out, open(os.path.join(ROOT, "profiles", f"{tag}_ncu_traffic.json"), "w"), indent=1)
    open(os.path.join(ROOT, "profiles", f"{tag}_chunk_kernels_ncu_summary.txt"), "w").write(
        "ncu --set full --clock-control none, one launch of each chunk kernel inside bench.py's timed regions "
        "(scripts/gpu_prof_r2.sh); under ncu every launch is serialised and cold\n" + "\n".join(text) + "\n")
    print(json.dumps(out))


if __name__ == "__main__":
    main()
