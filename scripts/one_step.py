"""One training step of a native engine inside an NVTX range named "capture" (for `ncu --nvtx --nvtx-include capture/`),
after warm-up steps outside it.  Writes the op labels of the step, in launch order, next to the report so the rows of
`ncu --page raw` can be named (scripts/ncu_table.py).

    python scripts/one_step.py --variant ref|milesial|milesial-bilinear [--batch 4] [--labels gpurun_out/labels_ref.json]
"""
import argparse
import json
import os

os.environ.setdefault("DPT_WGRAD_STREAM", "0")  # ops are timed / captured one by one on one stream
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.ops import native  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--variant", default="ref")
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--height", type=int, default=640)
ap.add_argument("--width", type=int, default=960)
ap.add_argument("--warmup", type=int, default=2)
ap.add_argument("--labels", default=None)
a = ap.parse_args()
torch.manual_seed(0)
B, H, W = a.batch, a.height, a.width
if a.variant == "ref":
    from distributedpytorch_b200.engine import NativeUNetEngine

    eng = NativeUNetEngine(UNet(), B, H, W, "cuda:0")
    eng.load_batch(torch.rand(B, 3, H, W, device="cuda"), (torch.rand(B, H, W, device="cuda") > 0.5).float())
else:
    from distributedpytorch_b200.engine.milesial import NativeMilesialEngine

    eng = NativeMilesialEngine(UNet(variant="milesial", n_classes=2, bilinear=a.variant.endswith("bilinear")), B, H, W, "cuda:0")
    eng.load_batch(torch.rand(B, 3, H, W, device="cuda"), (torch.rand(B, H, W, device="cuda") > 0.5).to(torch.uint8))
for _ in range(a.warmup):
    eng.train_step()
torch.cuda.synchronize()
# per-op launch counts, measured (an op may launch several kernels)
seq = []
for k in range(eng.step_end):
    before = native.launch_count()
    eng.fwd.run(k, k + 1)
    seq += [eng.fwd.labels[k]] * (native.launch_count() - before)
torch.cuda.synchronize()
if a.labels:
    json.dump(seq, open(a.labels, "w"))
torch.cuda.nvtx.range_push("capture")
eng.train_step()
torch.cuda.synchronize()
torch.cuda.nvtx.range_pop()
print(f"one step of {a.variant}: {len(seq)} kernel launches, loss {float(eng.loss_state[4]):.4f}")
