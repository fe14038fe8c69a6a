"""Per-source-line totals (stall samples, warp instructions, thread efficiency) from
`ncu -i X.ncu-rep --page source --print-source cuda,sass --csv`.  Usage: ncu_lines.py file.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
hdr = None; cur = None; agg = {}; src = {}
fname = None
for r in rows:
    if not r: continue
    if r[0] == "File Name": fname = r[1].split("/")[-1]; continue
    if r[0] == "Line No": hdr = r; continue
    if hdr is None or len(r) < len(hdr): continue
    if r[0] != "":
        cur = (fname, int(r[0])); src[cur] = r[1]
        continue
    i_s = hdr.index("# Samples"); i_i = hdr.index("Instructions Executed"); i_t = hdr.index("Thread Instructions Executed")
    i_l = hdr.index("stall_long_sb"); i_w = hdr.index("stall_wait"); i_sh = hdr.index("stall_short_sb")
    a = agg.setdefault(cur, [0, 0, 0, 0, 0, 0])
    for k, i in enumerate((i_s, i_i, i_t, i_l, i_w, i_sh)):
        try: a[k] += int(r[i])
        except ValueError: pass
ts = sum(a[0] for a in agg.values()); ti = sum(a[1] for a in agg.values())
print(f"total samples {ts}  warp-instructions {ti}")
print("file:line  samples%  inst%  thr/inst  long_sb  wait  short_sb | source")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{k[0]}:{k[1]:5d} {100*a[0]/ts:6.2f} {100*a[1]/max(ti,1):6.2f} {a[2]/max(a[1],1):5.1f} {a[3]:7d} {a[4]:6d} {a[5]:6d} | {src[k].strip()[:110]}")
