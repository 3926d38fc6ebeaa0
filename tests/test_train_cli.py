"""End-to-end plumbing on CPU (BASELINE.json config 1): the train.py command line in every mode on a tiny synthetic
dataset, checkpoint format, loss-curve pickles, logs, evaluate.py."""
import os
import subprocess
import sys

import pandas as pd
import pytest

from proc_util import run_group
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ENV = dict(os.environ, DPT_IMAGE_SIZE="48x32", DPT_SYNTHETIC="24", OMP_NUM_THREADS="2", CUDA_VISIBLE_DEVICES="",
           PYTHONPATH=ROOT)


def _run(args, cwd, timeout=900, launcher=None):
    cmd = (launcher or [sys.executable]) + [os.path.join(ROOT, "train.py"), "-v", "25"] + args
    r = run_group(cmd, timeout=timeout, env=ENV, cwd=cwd)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


def _check_outputs(cwd, mode, ckpt, prefixed):
    sd = torch.load(os.path.join(cwd, "checkpoints", f"{ckpt}.pth"), weights_only=True)
    assert len(sd) == 46
    key = ("module." if prefixed else "") + "encoder.conv1.conv_block.0.weight"
    assert key in sd and sd[key].dtype == torch.float32 and tuple(sd[key].shape) == (32, 3, 3, 3)
    tr = pd.read_pickle(os.path.join(cwd, "loss", mode, "train_loss.pkl"))
    va = pd.read_pickle(os.path.join(cwd, "loss", mode, "val_loss.pkl"))
    assert list(tr.columns) == ["Step", "Time", "Loss"] and list(va.columns) == ["Step", "Time", "Loss"]
    assert len(va) == 2 and va["Loss"].notna().all()
    log = open(os.path.join(cwd, "logs", f"{mode if mode != 'singleGPU' else 'singleGPU'}.log")).read()
    assert "UNet for Carvana Image Masking" in log


def test_single_gpu_mode_and_evaluate_cli(tmp_path):
    cwd = str(tmp_path)
    _run(["-e", "2", "-b", "2", "--lr", "1e-3"], cwd)
    _check_outputs(cwd, "singleGPU", "singleGPU", prefixed=False)
    r = run_group([sys.executable, os.path.join(ROOT, "evaluate.py"), "-c", "singleGPU", "-b", "2", "-v", "25"], timeout=600, env=ENV, cwd=cwd)
    assert r.returncode == 0 and "validation loss" in r.stdout, r.stdout + r.stderr[-2000:]
    # resume from the checkpoint through -c (the reference's backslash path cannot do this on Linux)
    _run(["-e", "1", "-b", "2", "-c", "singleGPU"], cwd)


def test_dp_mode(tmp_path):
    cwd = str(tmp_path)
    _run(["-t", "DP", "-e", "2", "-b", "4"], cwd)
    _check_outputs(cwd, "DP", "DP", prefixed=True)


def test_ddp_mode_two_ranks(tmp_path):
    cwd = str(tmp_path)
    launcher = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nnodes=1",
                "--nproc_per_node=2"]
    _run(["-t", "DDP", "-e", "2", "-b", "2"], cwd, launcher=launcher)
    _check_outputs(cwd, "DDP", "DDP", prefixed=True)


def test_mp_mode_two_stages(tmp_path):
    cwd = str(tmp_path)
    _run(["-t", "MP", "-e", "2", "-b", "4"], cwd)
    _check_outputs(cwd, "MP", "MP", prefixed=False)


def test_milesial_two_class_variant_on_cpu(tmp_path):
    cwd = str(tmp_path)
    _run(["-e", "2", "-b", "2", "--variant", "milesial", "--classes", "2"], cwd)
    sd = torch.load(os.path.join(cwd, "checkpoints", "singleGPU.pth"), weights_only=True)
    assert "inc.double_conv.0.weight" in sd and "outc.conv.weight" in sd and sd["outc.conv.weight"].shape[0] == 2
