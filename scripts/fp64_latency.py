#!/usr/bin/env python
"""FP64 dependent-issue latency of one warp (DFMA and DMMA m8n8k4), cycles per instruction for 1/2/4/8 independent chains."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import __graft_entry__ as ge
pkg = ge.load_package(); pkg.load_library()
ctx = pkg.Context(0)
for mma in (0, 1):
    print("DMMA" if mma else "DFMA", {ilp: round(ctx.fp64_latency(mma, ilp), 2) for ilp in (1, 2, 4, 8)})
