"""Pipeline timeline of the backward from an NFI_WS_TRACE dump (<file>.bwd.N, see nfi_field_bwd_ws.cuh)."""
import sys
import numpy as np
d = np.loadtxt(sys.argv[1], dtype=np.int64).reshape(5, -1, 6)[:, :, 2:]
t0 = d[d > 0].min()
lo, hi = int(sys.argv[2]) if len(sys.argv) > 2 else 40, int(sys.argv[3]) if len(sys.argv) > 3 else 52
hdr = ["G[meta ph1 aemp gath]", "D[start dfok read done]", "M[L0 L1 dH dF]", "E[acc0 epi0 acc1 heads]", "E[acc2 epi2 dfrd scat]"]
print("      " + " | ".join(hdr))
for t in range(lo, hi):
    print("t%3d" % t, end="")
    for r in (0, 1, 4, 2, 3):
        e = d[r, t][:4]
        print(" | " + " ".join("%6d" % (x - t0) if x > 0 else "     -" for x in e), end="")
    print()
for r, name in ((0, "gather start"), (4, "L0 issue")):
    x = d[r, :, 0]; x = x[x > 0]
    print(name, "mean period (cycles)", np.diff(np.sort(x)).mean())
g = d[0]; ok = (g[:, 3] > 0) & (g[:, 2] > 0)
print("phase1 %.0f  a_empty wait %.0f  gather %.0f" % ((g[ok, 1] - g[ok, 0]).mean(), (g[ok, 2] - g[ok, 1]).mean(), (g[ok, 3] - g[ok, 2]).mean()))
s = d[1]; ok = (s[:, 3] > 0) & (s[:, 1] > 0)
print("d-point: df wait %.0f  read %.0f  d-point %.0f" % ((s[ok, 1] - s[ok, 0]).mean(), (s[ok, 2] - s[ok, 1]).mean(), (s[ok, 3] - s[ok, 2]).mean()))
e = d[2]; ok = (e[:, 3] > 0) & (e[:, 0] > 0)
print("E: epi0 %.0f  wait L1 %.0f  heads %.0f" % ((e[ok, 1] - e[ok, 0]).mean(), (e[ok, 2] - e[ok, 1]).mean(), (e[ok, 3] - e[ok, 2]).mean()))
e3 = d[3]; ok = ok & (e3[:, 1] > 0)
ok = ok & (e3[:, 3] > 0)
print("E: wait dH %.0f  epi2 %.0f  wait dF + read %.0f  scatter %.0f" % ((e3[ok, 0] - e[ok, 3]).mean(), (e3[ok, 1] - e3[ok, 0]).mean(), (e3[ok, 2] - e3[ok, 1]).mean(), (e3[ok, 3] - e3[ok, 2]).mean()))
nxt = e[2:, 0] - e3[:-2, 3]; okn = (e[2:, 0] > 0) & (e3[:-2, 3] > 0)
print("E: scatter done -> next tile acc0 %.0f" % nxt[okn].mean())
m = d[4]; ok = (m[:, 3] > 0) & (m[:, 0] > 0)
print("MMA: L0->L1 %.0f  L1->dH %.0f  dH->dF %.0f" % ((m[ok, 1] - m[ok, 0]).mean(), (m[ok, 2] - m[ok, 1]).mean(), (m[ok, 3] - m[ok, 2]).mean()))
