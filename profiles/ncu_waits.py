"""Per mbarrier wait site (SASS TRYWAIT) of an ncu source CSV: stall samples + executions, with the
nearest enclosing line of the kernel's own source file, so waits can be attributed to a role."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1]))); main = sys.argv[2] if len(sys.argv) > 2 else "nfi_field_fwd_ws.cuh"
ins = []
f = None; line = None
for r in rows:
    if not r: continue
    if r[0] == "File Path": f = r[1].split('/')[-1]; continue
    if r[0] in ("Function Name", "Line No"): continue
    if r[0] != '': line = int(r[0]); continue
    try: ins.append((int(r[2], 16), f, line, r[3].strip(), int(r[6] or 0), int(r[7] or 0)))
    except Exception: pass
ins.sort()
tot = sum(i[4] for i in ins)
last = None
seen = set()
for k, (a, f, l, s, n, ex) in enumerate(ins):
    if f == main: last = l
    if "TRYWAIT" in s and a not in seen:
        seen.add(a)
        win = sum(i[4] for i in ins[max(0, k - 2):k + 6] if i[0] != a or True)
        print("%6.2f%% of samples  polls %9d  after %s:%s   %s" % (100.0 * win / tot, ex, main, last, s[:60]))
