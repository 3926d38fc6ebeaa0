"""Top stalled SASS instructions from `ncu -i X.ncu-rep --page source --csv` output (file, optionally .gz).
    python scripts/ncu_hot.py gpurun_out/ncu_src_ref_fwd_mid_2.csv.gz [N]"""
import csv
import gzip
import sys

path = sys.argv[1]
f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
rows = list(csv.reader(f))
hdr_i = next(i for i, r in enumerate(rows) if r and r[0] == "Address")
hdr = rows[hdr_i]
ix = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[hdr_i + 1:] if len(r) == len(hdr) and r[0].startswith("0x")]
stall_cols = [h for h in hdr if h.startswith("stall_")]


def num(x):
    try:
        return float(x)
    except ValueError:
        return 0.0


samp = "# Samples" if "# Samples" in ix else "Warp Stall Sampling (All Samples)"
tot = sum(num(r[ix[samp]]) for r in body) or 1.0
top = sorted(body, key=lambda r: -num(r[ix[samp]]))[: int(sys.argv[2]) if len(sys.argv) > 2 else 25]
print(rows[0][1] if len(rows[0]) > 1 else "", "| total samples", tot, "| instructions", len(body))
for r in top:
    s = num(r[ix[samp]])
    st = sorted(((num(r[ix[c]]), c) for c in stall_cols), reverse=True)[:3]
    print(f"{r[ix['Address']][-5:]} {s:8.0f} {100 * s / tot:5.1f}%  {r[ix['Source']].strip()[:64]:64s} " +
          " ".join(f"{c[6:]}={v:.0f}" for v, c in st if v))
