"""Top stall lines (with dominant reasons) of an `ncu --page source --csv --print-source cuda,sass` dump."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); topn = int(sys.argv[2]) if len(sys.argv) > 2 else 22
i = 0; out = []
while i < len(rows):
    if rows[i] and rows[i][0] == "File Path":
        f = rows[i][1].split('/')[-1]; hdr = rows[i + 2]; j = i + 3
        si = hdr.index("# Samples")
        cols = [k for k, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
        while j < len(rows) and not (rows[j] and rows[j][0] == "File Path"):
            r = rows[j]
            if r and r[0] != '' and len(r) > si:
                try: n = float(r[si])
                except Exception: n = 0
                if n > 0:
                    st = sorted([(float(r[k] or 0), hdr[k]) for k in cols if k < len(r)], reverse=True)[:3]
                    out.append((n, f, r[0], r[1].strip()[:70], st))
            j += 1
        i = j
    else:
        i += 1
tot = sum(o[0] for o in out)
for n, f, l, src, st in sorted(out, reverse=True)[:topn]:
    print("%5.1f%% %-22s %4s %-70s %s" % (100 * n / tot, f, l, src, ", ".join("%s %.0f" % (b.replace('stall_', ''), a) for a, b in st)))
