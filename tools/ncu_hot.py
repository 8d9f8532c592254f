"""Summarise an `ncu --page source --csv` dump: top SASS instructions by stall samples, with stall reasons."""
import csv, sys
path = sys.argv[1]; topn = int(sys.argv[2]) if len(sys.argv) > 2 else 40
rows = list(csv.reader(open(path)))
hdr = rows[1]; data = rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
def num(x):
    try: return float(x)
    except: return 0.0
tot = sum(num(r[ix["# Samples"]]) for r in data)
print("total samples", tot, "instructions", len(data))
agg = {}
for r in data:
    for c in stall_cols: agg[c] = agg.get(c, 0) + num(r[ix[c]])
print("stall mix:", {k: round(v / tot, 3) for k, v in sorted(agg.items(), key=lambda kv: -kv[1]) if v / tot > 0.005})
order = sorted(range(len(data)), key=lambda i: -num(data[i][ix["# Samples"]]))[:topn]
for i in sorted(order):
    r = data[i]
    st = sorted(((num(r[ix[c]]), c[6:]) for c in stall_cols), reverse=True)[:3]
    print(f"{i:5d} {num(r[ix['# Samples']]):8.0f} {100*num(r[ix['# Samples']])/tot:5.1f}% exec={num(r[ix['Instructions Executed']]):10.0f} {r[ix['Source']][:90]:90s} {[(s, int(v)) for v, s in st if v > 0]}")
