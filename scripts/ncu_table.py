"""Markdown table of the key ncu metrics (one row per captured launch) from `--page raw --csv` files,
plus the per-launch DRAM bytes as JSON.  Usage: ncu_table.py out.md out.json side raw1.csv [raw2.csv ...]"""
import csv, json, sys
out_md, out_json, side = sys.argv[1], sys.argv[2], int(sys.argv[3])
keys = [('gpu__time_duration.sum', 'time'), ('smsp__issue_active.avg.pct_of_peak_sustained_active', 'issue active %'),
        ('sm__warps_active.avg.pct_of_peak_sustained_active', 'warps active %'), ('launch__registers_per_thread', 'regs'),
        ('smsp__thread_inst_executed_per_inst_executed.ratio', 'threads / inst'), ('smsp__inst_executed.sum', 'warp inst'),
        ('dram__bytes_read.sum', 'dram read'), ('dram__bytes_write.sum', 'dram write'),
        ('l1tex__t_sector_hit_rate.pct', 'L1 hit %'), ('lts__t_sector_hit_rate.pct', 'L2 hit %'),
        ('smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio', 'stall long_sb'),
        ('smsp__average_warps_issue_stalled_wait_per_issue_active.ratio', 'stall wait'),
        ('smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio', 'stall short_sb'),
        ('smsp__average_warps_issue_stalled_branch_resolving_per_issue_active.ratio', 'stall branch'),
        ('smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio', 'stall not_selected')]
scale = {'byte': 1, 'Kbyte': 1e3, 'Mbyte': 1e6, 'Gbyte': 1e9}
lines = ["| kernel | " + " | ".join(k for _, k in keys) + " |", "|---|" + "---|" * len(keys)]
traffic = {}
for fn in sys.argv[4:]:
    rows = list(csv.reader(open(fn)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        rec = dict(zip(hdr, r)); un = dict(zip(hdr, units))
        name = rec['Kernel Name'].split('(')[0].replace('void ', '')
        vals = []
        for k, _ in keys:
            v = rec.get(k, '')
            try:
                f = float(v.replace(',', ''))
                v = ("%.3g" % f) if un[k] not in ('inst',) or k.startswith('smsp__average') else "%.3e" % f
            except ValueError:
                pass
            u = un.get(k, '')
            vals.append((v + (' ' + u if u in ('ms', 'us', 'Gbyte', 'Mbyte', 'Kbyte') else '')) if v else '-')
        lines.append("| %s | " % name + " | ".join(vals) + " |")
        b = sum(float(rec[k].replace(',', '')) * scale[un[k]] for k in ('dram__bytes_read.sum', 'dram__bytes_write.sum'))
        traffic.setdefault(name, []).append(b)
open(out_md, 'w').write("\n".join(lines) + "\n")
first = lambda n: (traffic.get(n) or [None])[0]
json.dump({"side": side,
           "command": "ncu --set full --clock-control none -k regex:<kernel> python bench.py --steps 1 --warmup 0 --no-cpu",
           "dram_bytes_per_launch": {"k_force": first('k_force<0>'), "k_density<0>": first('k_density<0>'),
                                     "k_density<1>": first('k_density<1>')}, "all": traffic}, open(out_json, 'w'), indent=1)
print("\n".join(lines))
