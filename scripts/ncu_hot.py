"""Top stalled SASS instructions from `ncu -i X.ncu-rep --page source --csv` output (stdin or file)."""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
hdr = rows[1]
ix = {h: i for i, h in enumerate(hdr)}
body = rows[2:]
stall_cols = [h for h in hdr if h.startswith("stall_") and "Not Issued" not in h]
def num(x):
    try: return float(x)
    except: return 0.0
tot = sum(num(r[ix["# Samples"]]) for r in body)
top = sorted(body, key=lambda r: -num(r[ix["# Samples"]]))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
print("total samples", tot)
for r in top:
    s = num(r[ix["# Samples"]])
    st = sorted(((num(r[ix[c]]), c) for c in stall_cols), reverse=True)[:3]
    print(f"{r[ix['Address']][-5:]} {s:8.0f} {100*s/tot:5.1f}%  {r[ix['Source']][:70]:70s} " + " ".join(f"{c[6:]}={v:.0f}" for v, c in st if v))
