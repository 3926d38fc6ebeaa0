"""Multi-GPU native engines on the box (skipped with fewer GPUs): single-process DataParallel and the pipeline."""
import json
import os
import subprocess
import sys

import pytest

from proc_util import run_group
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_native_dataparallel_matches_single_gpu():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = run_group([sys.executable, os.path.join(ROOT, "tests", "dp_check.py")], timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


def test_native_dataparallel_batchnorm_family():
    """North-star (BatchNorm) UNet on NativeDP: per-replica statistics like torch DataParallel, whole-batch loss."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    r = run_group([sys.executable, os.path.join(ROOT, "tests", "dp_check.py"), "milesial"], timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]


@pytest.mark.parametrize("variant", ["ref", "milesial"])
def test_native_dataparallel_multimem_arenas(variant):
    """One process, several GPUs: the replicas' arenas on VMM allocations under one multicast object, the fused update
    reducing with multimem.ld_reduce and broadcasting with multimem.st (same checks as the peer-access algorithms)."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    from distributedpytorch_b200.parallel.comm import multicast_supported

    if not all(multicast_supported(torch.device("cuda", d)) for d in range(torch.cuda.device_count())):
        pytest.skip("no NVSwitch multicast on this box")
    args = ["--algo=multimem"] if variant == "ref" else ["milesial", "--algo=multimem"]
    r = run_group([sys.executable, os.path.join(ROOT, "tests", "dp_check.py")] + args, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert '"algo": "multimem"' in r.stdout


@pytest.mark.parametrize("world,schedule,micro,port", [(2, "gpipe", 2, 29731), (2, "1f1b", 4, 29732), (4, "1f1b", 4, 29733)])
def test_native_pipeline_batchnorm_family(world, schedule, micro, port):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, DPT_SCHEDULE=schedule, DPT_MICRO=str(micro), DPT_BACKEND="nccl", DPT_VARIANT="milesial")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "pipe_worker.py")]
    r = run_group(cmd, timeout=300, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    res = [json.loads(l.split("PIPERESULT ", 1)[1]) for l in r.stdout.splitlines() if "PIPERESULT " in l]
    assert len(res) == world
    for x in res:
        # bf16 through 18 BatchNorm layers vs the fp32 oracle: Adam moves every weight by ~lr per step whatever the
        # gradient's size, so single weights may differ by a few lr; the mean must stay far below one step
        assert x["mean_param_diff"] < 1e-3 and x["max_param_diff"] < 1e-2, x
        for a, b in zip(x["losses"], x["losses_ref"]):
            assert abs(a - b) < 0.05 * abs(b) + 1e-2, x
        if x["losses"]:
            assert x["losses"][-1] < x["losses"][0], x


@pytest.mark.parametrize("world,schedule,micro,port", [(2, "gpipe", 2, 29721), (2, "1f1b", 2, 29722), (4, "gpipe", 4, 29723)])
def test_native_pipeline_tracks_unpipelined_fp32(world, schedule, micro, port):
    if torch.cuda.device_count() < world:
        pytest.skip(f"needs {world} GPUs")
    env = dict(os.environ, DPT_SCHEDULE=schedule, DPT_MICRO=str(micro), DPT_BACKEND="nccl")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "pipe_worker.py")]
    r = run_group(cmd, timeout=240, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    res = [json.loads(l.split("PIPERESULT ", 1)[1]) for l in r.stdout.splitlines() if "PIPERESULT " in l]
    assert len(res) == world
    for x in res:
        assert x["max_param_diff"] < 5e-3, x  # bf16 kernels vs fp32 oracle, 3 Adam steps at lr 1e-3
        for a, b in zip(x["losses"], x["losses_ref"]):
            assert abs(a - b) < 0.05 * abs(b) + 5e-3, x
