"""Native DDP (symmetric arenas + fused all-reduce/Adam/all-gather kernel) on 2 GPUs vs hand-rolled DDP
semantics on fp32 torch ops.  Skipped with fewer than 2 GPUs."""
import json
import os
import subprocess
import sys

import pytest

from proc_util import run_group
import torch

pytestmark = [pytest.mark.gpu, pytest.mark.multigpu]
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.mark.parametrize("algo", ["two_shot", "one_shot", "bucketed_overlap", "multimem"])
def test_fused_ddp_two_ranks(algo):
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    env = dict(os.environ, DPT_ALGO=algo)
    if algo == "bucketed_overlap":  # three buckets launched on the communication stream during backward
        env.update(DPT_ALGO="two_shot", DPT_DDP_OVERLAP="1")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29671", os.path.join(ROOT, "tests", "ddp_worker.py")]
    r = run_group(cmd, timeout=240, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    if "DDPSKIP" in r.stdout:
        pytest.skip("NVSwitch multicast (NVLS) is not available on this box")
    results = [json.loads(l.split("DDPRESULT ", 1)[1]) for l in r.stdout.splitlines() if "DDPRESULT " in l]
    assert len(results) == 2, r.stdout[-2000:]
    for res in results:
        assert res["same_init"], "rank-0 parameter broadcast failed"
        assert res["identical_replicas"], "replicas diverged after the fused update"
        assert res["min_param_cos"] > 0.999, res
        assert res["mean_update_cos"] > 0.9, res
        for a, b in zip(res["losses"], res["losses_ref"]):
            assert abs(a - b) < 0.05 * abs(b) + 5e-3, res


@pytest.mark.parametrize("algo", ["auto", "multimem"])
def test_fused_ddp_batchnorm_family_two_ranks(algo):
    """`multimem`: arenas AND the BatchNorm-buffer block on NVLS multicast objects; rank 0's parameters (constructor) and
    running statistics (every step) reach the other ranks through one multimem.st stream."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2", "--master-addr",
           "127.0.0.1", "--master-port", "29673" if algo == "auto" else "29674", os.path.join(ROOT, "tests", "ddp_worker_milesial.py")]
    r = run_group(cmd, timeout=240, env=dict(os.environ, DPT_ALGO=algo), cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    if "DDPSKIP" in r.stdout:
        pytest.skip("NVSwitch multicast (NVLS) is not available on this box")
    results = [json.loads(l.split("DDPRESULT ", 1)[1]) for l in r.stdout.splitlines() if "DDPRESULT " in l]
    assert len(results) == 2
    for res in results:
        assert res["identical_replicas"] and res["identical_bn_buffers"] and res["bn_moved"], res
        assert res["losses"][-1] < res["losses"][0], res


def test_fused_collective_barrier_stress_with_a_delayed_rank():
    """2000 back-to-back fused reduce-scatter + Adam + all-gather kernels per algorithm with a randomly chosen rank delayed
    by up to ~1 ms before some of them: the epoch-flag barriers must neither hang (watchdog) nor let a fast rank read
    gradients the slow rank has not written; replicas compared bit for bit every 100 steps."""
    n = torch.cuda.device_count()
    if n < 2:
        pytest.skip("needs 2 GPUs")
    world = 8 if n >= 8 else 4 if n >= 4 else 2
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29677", os.path.join(ROOT, "scripts", "bench_collective.py"), "--mb", "4",
           "--stress", "2000"]
    r = run_group(cmd, timeout=600, env=dict(os.environ), cwd=ROOT)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    res = [json.loads(l.split("STRESS ", 1)[1]) for l in r.stdout.splitlines() if l.startswith("STRESS ")]
    assert len(res) >= 2 and all(x["replicas_identical_every_100"] for x in res), res
