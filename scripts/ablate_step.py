"""What each op kind costs INSIDE the back-to-back training step: the step is re-timed with that kind left out of the
launch program (DPT_DEBUG_SKIP_OPS), one fresh process per ablation.  Isolated per-op timings (profile_step.py) include
launch gaps and cold caches; the difference of two back-to-back steps does not.
Usage: python scripts/ablate_step.py [--batch 4] > profiles/step_ablation.txt"""
import argparse
import os

os.environ.setdefault("DPT_WGRAD_STREAM", "0")  # ops are timed / captured one by one on one stream
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
CHILD = r"""
import sys, torch
sys.path.insert(0, %r)
from distributedpytorch_b200.engine import NativeUNetEngine
from distributedpytorch_b200.models import UNet
B = %d
torch.manual_seed(0)
eng = NativeUNetEngine(UNet(), B, 640, 960, "cuda:0")
eng.load_batch(torch.rand(B, 3, 640, 960, device="cuda"), (torch.rand(B, 640, 960, device="cuda") > 0.5).float())
for _ in range(5):
    eng.train_step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(30):
    eng.train_step()
e1.record()
torch.cuda.synchronize()
print("MS", e0.elapsed_time(e1) / 30, eng.step_end)
"""

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4)
args = ap.parse_args()


def run(skip):
    env = dict(os.environ, DPT_DEBUG_SKIP_OPS=skip)
    r = subprocess.run([sys.executable, "-c", CHILD % (ROOT, args.batch)], env=env, capture_output=True, text=True, timeout=300)
    for line in r.stdout.splitlines():
        if line.startswith("MS "):
            _, ms, n = line.split()
            return float(ms), int(n)
    raise RuntimeError(r.stderr[-2000:])


full, n_full = run("")
print(f"# batch {args.batch} 640x960, back-to-back step {full:.3f} ms, {n_full} ops")
print(f"# {'ops left out':28s} {'step ms':>9s} {'cost ms':>9s} {'ops':>5s}")
for skip in ("bwd bgrad", "bwd wgrad", "bwd dgrad", "fwd ", "pool", "head", "prep_weights", "adam", "pack_input"):
    ms, n = run(skip)
    print(f"  {skip:28s} {ms:9.3f} {full - ms:9.3f} {n_full - n:5d}")
