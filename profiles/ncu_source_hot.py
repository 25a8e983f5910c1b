"""Aggregate an `ncu --page source --csv --print-source cuda,sass` dump: per kernel, per CUDA source
line: executed warp instructions and stall samples (top N)."""
import csv, sys, collections
path, topn = sys.argv[1], int(sys.argv[2]) if len(sys.argv) > 2 else 25
def F(x):
    try:
        return float(x)
    except Exception:
        return 0.0


rows = list(csv.reader(open(path)))
i = 0
while i < len(rows):
    if rows[i] and rows[i][0] == "File Path":
        fpath = rows[i][1]; func = rows[i + 1][1]; hdr = rows[i + 2]
        j = i + 3
        body = []
        while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
            body.append(rows[j]) if (rows[j] and rows[j][0] != '') else None; j += 1
        if "Instructions Executed" in hdr and "Line No" in hdr and hdr[0] == "Line No":
            ci = hdr.index("Instructions Executed"); si = hdr.index("# Samples"); li = 0; srci = 1
            tot = sum(F(r[ci]) for r in body if len(r) > ci)
            tots = sum(F(r[si]) for r in body if len(r) > si)
            print("=== %s :: %s  (warp instr %.3g, samples %d)" % (fpath.split("/")[-1], func[:60], tot, tots))
            agg = sorted(body, key=lambda r: -F(r[ci]) if len(r) > ci else 0)[:topn]
            for r in agg:
                print("%5s  inst %5.1f%%  samp %5.1f%%  %s" % (r[li], 100 * F(r[ci]) / max(tot, 1),
                                                              100 * F(r[si]) / max(tots, 1), r[srci].strip()[:110]))
        i = j
    else:
        i += 1
