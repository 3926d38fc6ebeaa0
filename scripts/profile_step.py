"""Per-op device time of one native training step (each op of the launch program timed with CUDA events, median of
several repetitions).  Usage: python scripts/profile_step.py [--batch 4] [--reps 5] > profiles/step_ops.txt"""
import argparse
import os

os.environ.setdefault("DPT_WGRAD_STREAM", "0")  # ops are timed / captured one by one on one stream
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.engine import NativeUNetEngine  # noqa: E402
from distributedpytorch_b200.models import UNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--reps", type=int, default=5)
ap.add_argument("--height", type=int, default=640)
ap.add_argument("--width", type=int, default=960)
ap.add_argument("--variant", type=str, default="ref", help="ref | milesial (north-star BatchNorm UNet, 2 classes)")
args = ap.parse_args()

torch.manual_seed(0)
if args.variant == "milesial":
    from distributedpytorch_b200.engine.milesial import NativeMilesialEngine

    eng = NativeMilesialEngine(UNet(variant="milesial", n_classes=2), args.batch, args.height, args.width, "cuda:0")
    eng.load_batch(torch.rand(args.batch, 3, args.height, args.width, device="cuda"),
                   (torch.rand(args.batch, args.height, args.width, device="cuda") > 0.5).long())
else:
    eng = NativeUNetEngine(UNet(), args.batch, args.height, args.width, "cuda:0")
    eng.load_batch(torch.rand(args.batch, 3, args.height, args.width, device="cuda"),
                   (torch.rand(args.batch, args.height, args.width, device="cuda") > 0.5).float())
for _ in range(3):
    eng.train_step()
torch.cuda.synchronize()
n = eng.step_end
times = [[] for _ in range(n)]
for _ in range(args.reps):
    for k in range(n):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        eng.fwd.run(k, k + 1)
        e1.record()
        torch.cuda.synchronize()
        times[k].append(e0.elapsed_time(e1))
med = [sorted(t)[len(t) // 2] for t in times]
total = sum(med)
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
torch.cuda.synchronize()
e0.record()
for _ in range(10):
    eng.train_step()
e1.record()
torch.cuda.synchronize()
print(f"# batch {args.batch} {args.height}x{args.width}: sum of isolated ops {total:.3f} ms; back-to-back step "
      f"{e0.elapsed_time(e1) / 10:.3f} ms; {n} ops")
groups = {}
for k in range(n):
    lbl = eng.fwd.labels[k]
    key = " ".join(lbl.split()[:2])
    groups[key] = groups.get(key, 0.0) + med[k]
print("# by kind:")
for key, v in sorted(groups.items(), key=lambda kv: -kv[1]):
    print(f"#   {key:28s} {v:8.3f} ms  {100 * v / total:5.1f}%")
# forward + backward ms per pipeline block: the cost table of the partitioner (models/unet.py MEASURED_MS)
blocks = []
for b in range(9):
    f0, f1 = eng.fwd_block_range[b]
    b0, b1 = eng.bwd_block_range[b]
    blocks.append(round(sum(med[f0:f1]) + sum(med[b0:b1]), 3))
print(f"# by block (fwd+bwd ms): {blocks}")
for k in range(n):
    print(f"{k:4d} {med[k]:8.3f} ms  {eng.fwd.labels[k]}")
