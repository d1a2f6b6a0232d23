"""Summarise an ncu source-page CSV (sass view): executed-instruction mix and stall samples of the hot instructions.
usage: ncu -i X.ncu-rep --page source --csv --print-source sass > f.csv; python tools/ncu_hot.py f.csv [top_n]"""
import csv, sys, collections
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
data = rows[2:]
tot = sum(int(r[ix["Instructions Executed"]]) for r in data)
samp = sum(int(r[ix["# Samples"]]) for r in data)
print("kernel:", rows[0][1][:90], "| warp-inst executed", tot, "| samples", samp)
mix = collections.Counter(); smix = collections.Counter()
for r in data:
    op = r[ix["Source"]].split()
    op = [o for o in op if not o.startswith("@")][0].split(".")[0] if op else "?"
    mix[op] += int(r[ix["Instructions Executed"]]); smix[op] += int(r[ix["# Samples"]])
print("opcode: share of executed | share of samples")
for op, n in mix.most_common(22):
    print(f"  {op:10s} {100*n/tot:6.2f}%  {100*smix[op]/max(samp,1):6.2f}%")
stalls = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
agg = collections.Counter()
for r in data:
    for s in stalls:
        agg[s] += int(r[ix[s]] or 0)
print("stall samples:", ", ".join(f"{k[6:]} {100*v/max(samp,1):.1f}%" for k, v in agg.most_common(8)))
top = sorted(data, key=lambda r: -int(r[ix["# Samples"]]))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
print("hottest instructions by samples:")
for r in top:
    st = sorted(((int(r[ix[s]] or 0), s[6:]) for s in stalls), reverse=True)[:2]
    print(f"  {int(r[ix['# Samples']]):6d} {r[ix['Address']][-5:]} {r[ix['Source']].strip()[:70]:70s} {st}")
