import sys; sys.path.insert(0, "."); sys.path.insert(0, "tests")
import __graft_entry__ as ge
vx = ge.load_package()
ctx = vx.Context(0)
for m in (0, 1, 2):
    print("mode", m, ctx.fp64_peak(m))
