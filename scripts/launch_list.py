"""Prints the kernel launch list (name, ms, warp instructions) of an ncu csv written with
--metrics gpu__time_duration.sum,smsp__inst_executed.sum --csv --log-file FILE."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
h = [i for i, r in enumerate(rows) if r and r[0] == "ID"][0]
hdr = rows[h]
ik, im, iv = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value")
cur = {}
for r in rows[h + 1:]:
    if len(r) <= iv:
        continue
    e = cur.setdefault(r[0], [r[ik][:48], 0.0, 0.0])
    v = float(r[iv].replace(",", ""))
    if "time" in r[im]:
        e[1] = v
    else:
        e[2] = v
tot = sum(v[1] for v in cur.values())
for k, v in cur.items():
    print("%4s %-48s %9.3f ms %5.1f%%  %.3e inst" % (k, v[0], v[1] / 1e6, 100 * v[1] / tot, v[2]))
print("total %.3f ms" % (tot / 1e6))
