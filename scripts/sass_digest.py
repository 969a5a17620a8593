"""Opcode histogram per kernel of libwkv6_b200.so (cuobjdump -sass): the tracked proof that the hot kernels are
tcgen05 (UTCHMMA) / TMEM (LDTM) / TMA (UTMALDG, UTMASTG) code.  Writes profiles/<tag>_sass_digest.txt."""
import collections
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
SO = os.path.join(ROOT, "paper_accurate_fast_cheap_b200", "libwkv6_b200.so")
KEY = ["UTCHMMA", "UTCBAR", "LDTM", "STTM", "UTMALDG", "UTMASTG", "UTMAPF", "STSM", "LDSM", "SYNCS", "MUFU", "HMUL2", "HFMA2", "FFMA2",
       "FFMA", "FMUL", "FADD", "F2FP", "STS", "LDS", "SHFL", "BAR", "R2UR"]


def main():
    tag = sys.argv[1] if len(sys.argv) > 1 else "r2"
    out = subprocess.run(["cuobjdump", "-sass", SO], capture_output=True, text=True).stdout
    cur, hist = None, collections.OrderedDict()
    for line in out.splitlines():
        m = re.search(r"Function : (\S+)", line)
        if m:
            cur = subprocess.run(["c++filt", m.group(1)], capture_output=True, text=True).stdout.strip() or m.group(1)
            hist[cur] = collections.Counter()
            continue
        m = re.match(r"\s+/\*[0-9a-f]+\*/\s+(.*?);", line)
        if cur and m:
            t = m.group(1).split()
            op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
            hist[cur][op] += 1
    lines = [f"cuobjdump -sass {os.path.relpath(SO, ROOT)}: SASS instructions per kernel, selected opcodes (static counts)"]
    for k, h in hist.items():
        n = sum(h.values())
        lines.append(f"== {k[:110]}\n   {n} instructions; " + " ".join(f"{o}:{h[o]}" for o in KEY if h[o]))
    path = os.path.join(ROOT, "profiles", f"{tag}_sass_digest.txt")
    open(path, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))


if __name__ == "__main__":
    main()
