#!/usr/bin/env python
"""Per-source-line summary of an ncu report (cuda,sass correlated view).
usage: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > f.csv; python tools/ncu_lines.py f.csv [topN]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur_file = None
hdr = None
lines = []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur_file = r[1].split('/')[-1]
    elif len(r) > 10 and r[0] == 'Line No':
        hdr = r
    elif len(r) > 10 and r[0] != '':
        d = dict(zip(hdr, r))
        def f(k):
            try: return float(d[k])
            except Exception: return 0.0
        lines.append((cur_file, r[0], r[1].strip()[:90], f('Instructions Executed'), f('Thread Instructions Executed'), f('# Samples'),
                      f('stall_long_sb'), f('stall_short_sb'), f('stall_wait'), f('stall_branch_resolving'), f('stall_math'), f('stall_not_selected'), f('stall_no_inst')))
tot = sum(l[3] for l in lines); tott = sum(l[4] for l in lines); tots = sum(l[5] for l in lines)
print(f"total warp-inst {tot:.3e} thread-inst {tott:.3e} avg threads {tott/max(tot,1):.2f} samples {tots:.0f}")
lines.sort(key=lambda l: -l[3])
print("file:line  inst%  avgthr  samp%  [long short wait br math notsel noinst]  source")
for l in lines[:top]:
    print(f"{l[0]}:{l[1]:>4} {100*l[3]/tot:5.1f} {l[4]/max(l[3],1):5.1f} {100*l[5]/max(tots,1):5.1f}  [{l[6]:.0f} {l[7]:.0f} {l[8]:.0f} {l[9]:.0f} {l[10]:.0f} {l[11]:.0f} {l[12]:.0f}]  {l[2]}")

# region summary (rt_device.cuh line ranges)
if len(sys.argv) > 3:
    import ast
    regions = ast.literal_eval(open(sys.argv[3]).read())
    agg = {}
    for l in lines:
        name = 'other:' + l[0]
        if l[0] == 'rt_device.cuh':
            for (a, b, n) in regions:
                if a <= int(l[1]) <= b: name = n; break
        a = agg.setdefault(name, [0, 0, 0])
        a[0] += l[3]; a[1] += l[4]; a[2] += l[5]
    print("\nregion  inst%  avgthr  samples%")
    for n, a in sorted(agg.items(), key=lambda x: -x[1][0]):
        print(f"{n:28s} {100*a[0]/tot:5.1f} {a[1]/max(a[0],1):5.1f} {100*a[2]/max(tots,1):5.1f}")
