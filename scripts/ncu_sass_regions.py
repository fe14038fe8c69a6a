"""Attributes every SASS instruction of one kernel (ncu source page, cuda,sass view) to the INNERMOST
source line it is listed under first, and sums stall samples / executed instructions per region of
kernels.cuh given as line ranges.  Usage: ncu_sass_regions.py file.csv name:lo-hi ..."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
regions = []
for a in sys.argv[2:]:
    nm, r = a.split(":"); lo, hi = r.split("-"); regions.append((nm, int(lo), int(hi)))
hdr = None; cur = None
seen = {}   # address -> (line, samples, inst, thread_inst)
for r in rows:
    if not r: continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        try: cur = int(r[0])
        except ValueError: cur = None
        continue
    addr = r[2]
    if addr in ("...", "-") or cur is None: continue
    try:
        s = int(r[hdr.index("# Samples")]); i = int(r[hdr.index("Instructions Executed")]); t = int(r[hdr.index("Thread Instructions Executed")])
    except ValueError: continue
    # keep the attribution to the deepest (largest line number inside helper range first seen) -> choose first seen
    if addr not in seen: seen[addr] = [set(), s, i, t]
    seen[addr][0].add(cur)
ts = sum(v[1] for v in seen.values()); ti = sum(v[2] for v in seen.values())
print("unique SASS instr %d, samples %d, warp-inst %d" % (len(seen), ts, ti))
agg = {nm: [0, 0, 0] for nm, _, _ in regions}; agg["other"] = [0, 0, 0]
for addr, (lines, s, i, t) in seen.items():
    hit = "other"
    for nm, lo, hi in regions:          # first region (in argument order) containing any of the lines wins
        if any(lo <= l <= hi for l in lines): hit = nm; break
    agg[hit][0] += s; agg[hit][1] += i; agg[hit][2] += t
for nm, (s, i, t) in agg.items():
    print("%-12s samples %5.1f%%  inst %5.1f%%  thr/inst %4.1f" % (nm, 100 * s / ts, 100 * i / max(ti, 1), t / max(i, 1)))
