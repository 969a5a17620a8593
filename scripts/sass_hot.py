"""Summarises `ncu -i rep --page source --csv` (SASS view): stall samples and issued instructions by opcode and the
hottest instructions.  Usage: ncu -i x.ncu-rep --page source --csv > x.csv; python scripts/sass_hot.py x.csv [kernel-substr]"""
import collections
import csv
import sys

rows = list(csv.reader(open(sys.argv[1])))
want = sys.argv[2] if len(sys.argv) > 2 else None
blocks, cur = [], None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "rows": []}
        blocks.append(cur)
    elif cur is not None and r:
        cur["rows"].append(r)
for blk in blocks:
    if want and want not in blk["name"]:
        continue
    hdr, data = blk["rows"][0], blk["rows"][1:]
    ia, isamp, iex = hdr.index("Source"), hdr.index("# Samples"), hdr.index("Instructions Executed")
    data = [r for r in data if len(r) > max(ia, isamp, iex) and r[isamp].isdigit()]
    tot = sum(int(r[isamp]) for r in data) or 1
    totex = sum(int(r[iex]) for r in data if r[iex].isdigit()) or 1
    print("==", blk["name"][:60], "samples", tot, "warp-instr", totex, "sass lines", len(data))
    byop, exop = collections.Counter(), collections.Counter()
    for r in data:
        t = r[ia].split()
        op = (t[1] if t[0].startswith("@") else t[0]).split(".")[0]
        byop[op] += int(r[isamp]); exop[op] += int(r[iex]) if r[iex].isdigit() else 0
    for op, n in byop.most_common(18):
        print(f"  {op:10s} samples {100*n/tot:5.1f}%   instr {100*exop[op]/totex:5.1f}%")
    for r in sorted(data, key=lambda r: -int(r[isamp]))[:12]:
        print("  ", r[isamp], r[iex], r[ia][:100])
    break
