"""Markdown table from an ncu report of one training step (scripts/capture_all.sh): one row per kernel launch, named by
the op label of the launch program, with the metrics the roofline argument needs.

    python scripts/ncu_table.py gpurun_out/ncu_raw_ref.csv gpurun_out/labels_ref.json > profiles/ncu_r2_step_ref.md
"""
import csv
import io
import json
import subprocess
import sys

rep, labels = sys.argv[1], json.load(open(sys.argv[2])) if len(sys.argv) > 2 else None
METRICS = {
    "gpu__time_duration.sum": "us",
    "TPC.TriageCompute.sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed": "tensor pipe %",
    "l1tex__data_pipe_tc_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed": "tc smem read %",
    "dram__bytes_read.sum": "dram rd MB",
    "dram__bytes_write.sum": "dram wr MB",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram %",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "L2 %",
    "lts__t_sector_hit_rate.pct": "L2 hit %",
    "sm__throughput.avg.pct_of_peak_sustained_elapsed": "SM %",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "occ %",
    "launch__registers_per_thread": "regs",
    "launch__grid_size": "grid",
    "launch__shared_mem_per_block_dynamic": "dyn smem KB",
}
if rep.endswith(".csv"):  # exported on the GPU box by scripts/ncu_export.py
    out = open(rep).read()
else:
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
rows = list(csv.reader(io.StringIO(out)))
hdr, units, body = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
cols = [(m, n) for m, n in METRICS.items() if m in ix]
tensor_cols = [h for h in hdr if "pipe_tensor" in h and "pct_of_peak_sustained_active" in h]


def val(r, m):
    try:
        v = float(r[ix[m]].replace(",", ""))
    except (ValueError, KeyError):
        return ""
    u = units[ix[m]]
    if m.startswith("gpu__time"):
        v = v / 1e3 if u in ("ns", "nsecond") else v * 1e3 if u in ("ms", "msecond") else v
        return f"{v:.1f}"
    if m.startswith("launch__shared"):
        return f"{v / 1024 if u == 'byte' else v:.0f}"
    if "bytes" in m:
        scale = {"byte": 1e-6, "Kbyte": 1e-3, "Mbyte": 1.0, "Gbyte": 1e3}.get(u, 1e-6)
        return f"{v * scale:.1f}"
    return f"{v:.1f}" if v != int(v) else str(int(v))


print(f"ncu --set full, one training step, batch 4 x 640 x 960 ({rep}); {len(body)} launches\n")
names = ["#", "op", "kernel"] + [n for _, n in cols]
print("| " + " | ".join(names) + " |")
print("|" + "---|" * len(names))
total = 0.0
for k, r in enumerate(body):
    kn = r[ix["Kernel Name"]]
    kn = kn.replace("dpt::", "").split("(")[0][:44]
    lab = labels[k] if labels and k < len(labels) else ""
    cells = [val(r, m) for m, _ in cols]
    try:
        total += float(cells[0])
    except ValueError:
        pass
    print("| " + " | ".join([str(k), lab, kn] + cells) + " |")
print(f"\nsum of kernel durations under ncu (serialised, cache-flushed): {total / 1e3:.3f} ms")
if tensor_cols:
    print("tensor-pipe metrics present in the report: " + ", ".join(tensor_cols[:6]))
