"""Print a pipeline timeline from an NFI_WS_TRACE dump (see nfi_field_fwd_ws.cuh)."""
import sys
import numpy as np
d = np.loadtxt(sys.argv[1], dtype=np.int64).reshape(5, -1, 6)[:, :, 2:]
t0 = d[d > 0].min()
lo, hi = int(sys.argv[2]) if len(sys.argv) > 2 else 40, int(sys.argv[3]) if len(sys.argv) > 3 else 56
names = {0: "P0 [meta_ok ph1 a_empty_ok gathered]", 1: "P1", 2: "E  [d0_ok epi0 d1_ok done]", 3: "MMA[a_full_ok L0 h_ok L1(prev)]"}
for t in range(lo, hi):
    print("tile %3d" % t, end="")
    for r in (0, 1, 3, 2):
        e = d[r, t]
        print(" | %s" % " ".join("%6d" % ((x - t0)) if x > 0 else "     -" for x in e), end="")
    print()
print("roles:", "  ".join(names[r] for r in (0, 1, 3, 2)))
# per-role period
for r in (3, 2):
    x = d[r, :, 0]; x = x[x > 0]
    print("role", r, "mean period (cycles)", np.diff(x).mean() if len(x) > 1 else None)
