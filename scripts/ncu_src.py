#!/usr/bin/env python
"""Summarises `ncu -i X.ncu-rep --page source --csv` (stdin or file): per kernel, executed instructions per warp and the
instructions that collect the most stall samples."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]) if len(sys.argv) > 1 else sys.stdin))
ntop = int(sys.argv[2]) if len(sys.argv) > 2 else 40
secs = []; cur = None
for r in rows:
    if r and r[0] == "Kernel Name":
        cur = {"name": r[1], "hdr": None, "data": []}; secs.append(cur)
    elif cur is not None and cur["hdr"] is None:
        cur["hdr"] = r
    elif cur is not None and len(r) == len(cur["hdr"]):
        cur["data"].append(r)
for s in secs:
    h = s["hdr"]; d = s["data"]
    iS = h.index("Source"); iW = h.index("Warp Stall Sampling (All Samples)"); iE = h.index("Instructions Executed")
    tot = sum(int(r[iE]) for r in d); tots = sum(int(r[iW]) for r in d)
    first = max(int(d[0][iE]), 1)
    print("== %s\n   static %d, executed %d warp-instr (%.0f per warp), %d samples" % (s["name"][:60], len(d), tot, tot / first, tots))
    top = sorted(enumerate(d), key=lambda t: -int(t[1][iW]))[:ntop]
    for i, r in sorted(top):
        print("   %5d %-72s %6s %9s" % (i, r[iS].strip()[:72], r[iW], r[iE]))
