"""Per-source-line summary of `ncu -i rep --page source --csv --print-source cuda,sass`: samples and warp-instructions.
Usage: python scripts/src_hot.py file.csv [topN]"""
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None
lines = {}
fname = ""
for r in rows:
    if r and r[0] == "File Path":
        fname = r[1].split("/")[-1]
    elif r and r[0] == "Line No":
        hdr = r
        isamp, iex = hdr.index("# Samples"), hdr.index("Instructions Executed")
    elif hdr and len(r) > iex and r[0].isdigit() and r[2] == "-":      # a source line row (address column is "-")
        key = (fname, int(r[0]))
        s, e = int(r[isamp] or 0), int(r[iex] or 0)
        if key in lines:
            lines[key] = (lines[key][0], lines[key][1] + s, lines[key][2] + e)
        else:
            lines[key] = (r[1], s, e)
    if len(lines) and r and r[0] == "File Path" and len(lines) > 0 and fname and False:
        pass
ts = sum(v[1] for v in lines.values()) or 1
te = sum(v[2] for v in lines.values()) or 1
print("total samples", ts, "warp-instr", te)
for (f, ln), (src, s, e) in sorted(lines.items(), key=lambda kv: -kv[1][1])[:top]:
    print(f"{f}:{ln:4d} samp {100*s/ts:5.1f}% instr {100*e/te:5.1f}%  {src.strip()[:110]}")
