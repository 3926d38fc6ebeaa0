"""Runs ON the GPU box after scripts/capture_all.sh: the .ncu-rep of a whole training step is too large to bring back
(180 MB with source), so export what is read later - the raw metric page of every launch (CSV) and the per-instruction
source page of the launches named below - and drop the report.

    python scripts/ncu_export.py gpurun_out/ncu_step_ref.ncu-rep gpurun_out/labels_ref.json ref
"""
import gzip
import json
import os
import subprocess
import sys

rep, labels_file, tag = sys.argv[1], sys.argv[2], sys.argv[3]
labels = json.load(open(labels_file))
out_dir = os.path.dirname(rep)
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True)
open(os.path.join(out_dir, f"ncu_raw_{tag}.csv"), "w").write(raw.stdout)
# source pages: first launch of each op whose label contains one of these
WANT = {
    "ref": ["fwd enc1.2", "fwd mid.2", "fwd dec4.0", "fwd deconv4", "bwd wgrad+bgrad encoder.conv1.conv_block.2",
            "bwd wgrad+bgrad mid.conv_block.2", "bwd dgrad decoder.conv4.conv_block.0", "bwd dgrad mid.conv_block.2",
            "bwd dgrad decoder.deconv4", "fwd head+loss", "bwd head", "bwd pool1", "fwd pool1", "adam", "prep_weights",
            "pack_input"],
    "milesial": ["fwd conv+stats inc.double_conv.3", "fwd conv+stats down4.maxpool_conv.1.double_conv.3",
                 "fwd bn apply+relu+pool inc", "fwd bn finalize inc.double_conv.4", "bwd bn up4.conv.double_conv.4",
                 "fwd outc+softmax", "bwd outc+softmax", "rmsprop"],
}.get(tag, [])
done = []
for w in WANT:
    idx = next((k for k, lab in enumerate(labels) if w in lab), None)
    if idx is None:
        continue
    name = w.replace(" ", "_").replace("+", "_").replace(".", "_")
    r = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--launch-skip", str(idx), "--launch-count", "1"],
                       capture_output=True, text=True)
    with gzip.open(os.path.join(out_dir, f"ncu_src_{tag}_{name}.csv.gz"), "wt") as f:
        f.write(r.stdout)
    done.append((w, idx))
print(f"exported raw page ({len(raw.stdout) >> 10} KiB) and {len(done)} source pages: {done}")
if os.path.getsize(rep) > 40 << 20:
    os.remove(rep)
