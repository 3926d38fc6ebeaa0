"""Forward / backward / update time of the native step, measured inside back-to-back steps with CUDA events on the step's
stream (the backward's second stream joins before the update, so the event after the backward range covers it).

    DPT_WGRAD_STREAM=0|1 python scripts/phase_times.py [--variant ref|milesial] [--steps 30]
"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--variant", default="ref")
ap.add_argument("--steps", type=int, default=30)
a = ap.parse_args()
B, H, W = 4, 640, 960
torch.manual_seed(0)
if a.variant == "ref":
    from distributedpytorch_b200.engine import NativeUNetEngine

    eng = NativeUNetEngine(UNet(), B, H, W, "cuda:0")
    eng.load_batch(torch.rand(B, 3, H, W, device="cuda"), (torch.rand(B, H, W, device="cuda") > 0.5).float())
else:
    from distributedpytorch_b200.engine.milesial import NativeMilesialEngine

    eng = NativeMilesialEngine(UNet(variant="milesial", n_classes=2), B, H, W, "cuda:0")
    eng.load_batch(torch.rand(B, 3, H, W, device="cuda"), (torch.rand(B, H, W, device="cuda") > 0.5).to(torch.uint8))
aux = getattr(eng, "aux_stream", None)
for _ in range(5):
    eng.train_step()
torch.cuda.synchronize()
marks = [[torch.cuda.Event(enable_timing=True) for _ in range(4)] for _ in range(a.steps)]
bounds = [0, eng.fwd_end, eng.bwd_end, eng.step_end]
if a.variant != "ref":
    eng._enter_train(None)
for ev in marks:
    ev[0].record()
    for k in range(3):
        eng.fwd.run(bounds[k], bounds[k + 1], None, aux)
        ev[k + 1].record()
torch.cuda.synchronize()
ms = [sum(ev[k].elapsed_time(ev[k + 1]) for ev in marks) / a.steps for k in range(3)]
print(json.dumps({"variant": a.variant, "two_stream_backward": bool(getattr(eng, "wgrad_concurrent", False)),
                  "forward_ms": round(ms[0], 3), "backward_ms": round(ms[1], 3), "update_ms": round(ms[2], 3),
                  "step_ms": round(sum(ms), 3), "steps": a.steps}))
