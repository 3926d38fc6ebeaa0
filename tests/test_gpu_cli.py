"""The train.py command line on the GPU box: every mode goes through the native engines (the library must be
loaded; there is no silent fallback), writes reference-format checkpoints, and the loss goes down."""
import os
import subprocess
import sys

import pandas as pd
import pytest

from proc_util import run_group
import torch

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _env():
    return dict(os.environ, DPT_IMAGE_SIZE="96x64", DPT_SYNTHETIC="40", PYTHONPATH=ROOT)


def _run(args, cwd, launcher=None, timeout=900):
    cmd = (launcher or [sys.executable]) + [os.path.join(ROOT, "train.py"), "-v", "20", "--lr", "1e-3"] + args
    r = run_group(cmd, timeout=timeout, env=_env(), cwd=cwd)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    return r


def _check(cwd, mode, prefixed):
    sd = torch.load(os.path.join(cwd, "checkpoints", f"{mode}.pth"), weights_only=True)
    key = ("module." if prefixed else "") + "decoder.deconv1.weight"
    assert len(sd) == 46 and tuple(sd[key].shape) == (512, 256, 2, 2)
    assert all(torch.isfinite(v).all() for v in sd.values())
    va = pd.read_pickle(os.path.join(cwd, "loss", mode, "val_loss.pkl"))
    assert va["Loss"].notna().all() and va["Loss"].iloc[-1] < va["Loss"].iloc[0] + 0.05
    log = open(os.path.join(cwd, "logs", f"{mode}.log")).read()
    return log


def test_cli_single_gpu_native(tmp_path):
    cwd = str(tmp_path)
    _run(["-e", "3", "-b", "4"], cwd)
    log = _check(cwd, "singleGPU", False)
    assert "native sm_100a engine" in log
    r = subprocess.run([sys.executable, os.path.join(ROOT, "evaluate.py"), "-c", "singleGPU", "-b", "4", "-v", "20"],
                       capture_output=True, text=True, timeout=600, env=_env(), cwd=cwd)
    assert r.returncode == 0 and "dice" in r.stdout, r.stdout + r.stderr[-2000:]


@pytest.mark.multigpu
def test_cli_dp_native(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cwd = str(tmp_path)
    _run(["-t", "DP", "-e", "3", "-b", "4"], cwd)
    assert "native" in _check(cwd, "DP", True)


@pytest.mark.multigpu
def test_cli_ddp_native(tmp_path):
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    cwd = str(tmp_path)
    launcher = [sys.executable, "-m", "torch.distributed.run", "--standalone", "--local-addr", "127.0.0.1", "--nnodes=1",
                "--nproc_per_node=2"]
    _run(["-t", "DDP", "-e", "3", "-b", "2"], cwd, launcher=launcher)
    assert "native fused" in _check(cwd, "DDP", True)


@pytest.mark.multigpu
def test_cli_mp_native(tmp_path):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cwd = str(tmp_path)
    _run(["-t", "MP", "-e", "3", "-b", "4"], cwd)
    _check(cwd, "MP", False)


def _check_milesial(cwd, mode, prefixed):
    sd = torch.load(os.path.join(cwd, "checkpoints", f"{mode}.pth"), weights_only=True)
    key = ("module." if prefixed else "") + "down4.maxpool_conv.1.double_conv.0.weight"
    assert tuple(sd[key].shape) == (1024, 512, 3, 3) and all(torch.isfinite(v).all() for v in sd.values() if v.is_floating_point())
    rm = sd[("module." if prefixed else "") + "up4.conv.double_conv.4.running_mean"]
    assert float(rm.abs().sum()) > 0, "BatchNorm running statistics were not trained / merged into the checkpoint"
    va = pd.read_pickle(os.path.join(cwd, "loss", mode, "val_loss.pkl"))
    assert va["Loss"].notna().all()
    return open(os.path.join(cwd, "logs", f"{mode}.log")).read()


@pytest.mark.multigpu
def test_cli_dp_native_batchnorm_variant(tmp_path):
    """`-t DP --variant milesial` runs on NativeDP (round 1 fell back to torch ops on one device)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cwd = str(tmp_path)
    _run(["-t", "DP", "-e", "2", "-b", "4", "--variant", "milesial"], cwd)
    log = _check_milesial(cwd, "DP", True)
    assert "native path, milesial variant" in log, log


@pytest.mark.multigpu
def test_cli_mp_native_batchnorm_variant(tmp_path):
    """`-t MP --variant milesial` runs on NativeStage (round 1 raised NotImplementedError)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cwd = str(tmp_path)
    _run(["-t", "MP", "-e", "2", "-b", "4", "--variant", "milesial", "--micro-batches", "2"], cwd)
    log = _check_milesial(cwd, "MP", False)
    assert "native stages (milesial variant)" in log, log


def test_cli_resume_native(tmp_path):
    cwd = str(tmp_path)
    _run(["-e", "1", "-b", "4"], cwd)
    _run(["-e", "2", "-b", "4", "--resume"], cwd)
    log = open(os.path.join(cwd, "logs", "singleGPU.log")).read()
    assert "Resumed singleGPU at epoch 1" in log
    va = pd.read_pickle(os.path.join(cwd, "loss", "singleGPU", "val_loss.pkl"))
    assert len(va) == 2
