"""Per-source-line hot spots of one kernel: ncu -i X.ncu-rep --page source --csv --print-source cuda,sass > f.csv;
python tools/ncu_lines.py f.csv [top]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
top = int(sys.argv[2]) if len(sys.argv) > 2 else 40
cur, hdr, agg = None, None, []
for r in rows:
    if len(r) == 2 and r[0] == 'File Path':
        cur = r[1].split('/')[-1]; continue
    if r and r[0] == 'Line No':
        hdr = r; continue
    if hdr and len(r) == len(hdr) and r[0].isdigit():
        d = dict(zip(hdr, r))
        agg.append((int(d['Instructions Executed']), int(d['# Samples']), int(d['Thread Instructions Executed']), cur, r[0], r[1].strip()[:100]))
ti = sum(a[0] for a in agg); ts = sum(a[1] for a in agg)
print(f"warp-instructions {ti}  samples {ts}")
byfile = {}
for a in agg:
    f = byfile.setdefault(a[3], [0, 0]); f[0] += a[0]; f[1] += a[1]
for f, v in sorted(byfile.items(), key=lambda kv: -kv[1][0]):
    print(f"  {f:28s} inst {100 * v[0] / ti:5.1f}%  samples {100 * v[1] / ts:5.1f}%")
for a in sorted(agg, reverse=True)[:top]:
    print(f"{100 * a[0] / ti:5.1f}% smp {100 * a[1] / ts:5.1f}% thr/inst {a[2] / max(a[0], 1):4.1f} {a[3]}:{a[4]:>4s} {a[5]}")
