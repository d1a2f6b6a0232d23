"""Summarise an ncu launch list (--metrics gpu__time_duration.sum --csv): python tools/launch_summary.py file.csv [min_us]"""
import csv, collections, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = [i for i, r in enumerate(rows) if r and r[0] == 'ID'][0]
h = rows[hdr]; ix = {n: i for i, n in enumerate(h)}
agg = collections.OrderedDict()
for r in rows[hdr + 1:]:
    if len(r) < len(h): continue
    k = r[ix['Kernel Name']].split('(')[0][-60:]
    v = float(r[ix['Metric Value']].replace(',', '')); u = r[ix['Metric Unit']]
    v = v / 1e3 if u in ('ns', 'nsecond') else v
    agg.setdefault(k, []).append(v)
tot = sum(sum(v) for v in agg.values())
mn = float(sys.argv[2]) if len(sys.argv) > 2 else 0
for k, v in agg.items():
    if sum(v) / len(v) >= mn:
        print(f"{k:60s} n={len(v):4d} mean={sum(v)/len(v):10.1f} us  share={100*sum(v)/tot:5.1f}%")
