"""--resume: kill-and-resume equals an uninterrupted run (CPU plumbing path).  The reference can only reload weights
(train.py:42-43); `<mode>.resume.pth` adds optimizer moments, scheduler, epoch / step position, curves and RNG state."""
import os
import sys

import pandas as pd
import torch

from proc_util import run_group

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, DPT_IMAGE_SIZE="48x32", DPT_SYNTHETIC="16", OMP_NUM_THREADS="1", CUDA_VISIBLE_DEVICES="",
           PYTHONPATH=ROOT)


def _train(args, cwd):
    cmd = [sys.executable, os.path.join(ROOT, "train.py"), "-v", "25", "-b", "2", "--lr", "1e-3"] + args
    r = run_group(cmd, timeout=900, env=ENV, cwd=cwd)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


def test_resume_continues_where_the_interrupted_run_stopped(tmp_path):
    full, cut = str(tmp_path / "full"), str(tmp_path / "cut")
    os.makedirs(full), os.makedirs(cut)
    _train(["-e", "3"], full)                       # uninterrupted: 3 epochs
    _train(["-e", "2"], cut)                        # "killed" after epoch 2 ...
    state = torch.load(os.path.join(cut, "checkpoints", "singleGPU.resume.pth"), weights_only=False)
    assert state["epoch"] == 2 and state["global_step"] == 12 and state["optimizer"] is not None
    _train(["-e", "3", "--resume"], cut)            # ... and resumed for the third
    log = open(os.path.join(cut, "logs", "singleGPU.log")).read()
    assert "Resumed singleGPU at epoch 2, step 12" in log
    a = torch.load(os.path.join(full, "checkpoints", "singleGPU.pth"), weights_only=True)
    b = torch.load(os.path.join(cut, "checkpoints", "singleGPU.pth"), weights_only=True)
    worst = max(float((a[k] - b[k]).abs().max()) for k in a)
    assert worst < 1e-5, f"resumed weights differ from the uninterrupted run by {worst}"
    va, vb = (pd.read_pickle(os.path.join(d, "loss", "singleGPU", "val_loss.pkl")) for d in (full, cut))
    assert len(va) == len(vb) == 3 and list(va["Step"]) == list(vb["Step"])
    assert abs(va["Loss"].iloc[-1] - vb["Loss"].iloc[-1]) < 1e-5


def test_missing_dataset_raises_instead_of_training_on_synthetic_data(tmp_path):
    env = {k: v for k, v in ENV.items() if k != "DPT_SYNTHETIC"}
    r = run_group([sys.executable, os.path.join(ROOT, "train.py"), "-e", "1"], timeout=300, env=env, cwd=str(tmp_path))
    assert r.returncode != 0 and "DPT_SYNTHETIC" in (r.stdout + r.stderr)
    assert not os.path.exists(os.path.join(str(tmp_path), "checkpoints", "singleGPU.pth"))
