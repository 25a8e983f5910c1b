"""Kernel sequence of ONE resident-input step from an `ncu --metrics gpu__time_duration.sum --csv` launch
list of bench.py (steps are delimited by the NCHW -> channels-last plane transpose)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hi = [i for i, r in enumerate(rows) if 'Kernel Name' in r][0]
h = rows[hi]; kn = h.index('Kernel Name'); mv = h.index('Metric Value')
seq = [(r[kn], float(r[mv].replace(',', '')) / 1e3) for r in rows[hi + 1:] if len(r) > mv]
starts = [i for i, (k, v) in enumerate(seq) if 'plane_transpose' in k and '<1>' in k]
n = int(sys.argv[2]) if len(sys.argv) > 2 else 3
a, b = starts[n], starts[n + 1]
tot = sum(v for _, v in seq[a:b])
for k, v in seq[a:b]:
    if v >= 15: print("%9.1f us %5.1f%%  %s" % (v, 100 * v / tot, k[:100]))
print("total %.1f us over %d launches (%.1f us in launches < 15 us)" % (tot, b - a, sum(v for _, v in seq[a:b] if v < 15)))
