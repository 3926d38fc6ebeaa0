"""Pipeline engine on CPU with gloo (world 2 and 4): partitioner, direct skip routing, GPipe / 1F1B schedules.
Pipelined training must equal un-pipelined training (exact for this BatchNorm-free model)."""
import json
import os
import subprocess
import sys

import pytest

from proc_util import run_group

from distributedpytorch_b200.parallel import pipe as P

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_partitioner_and_flows():
    costs = [12.39, 16.99, 16.99, 16.99, 16.99, 36.5, 36.5, 36.5, 36.54]
    assert P.partition(costs, 1) == [0] * 9
    two = P.partition(costs, 2)
    assert two == sorted(two) and set(two) == {0, 1}
    loads = [sum(c for c, s in zip(costs, two) if s == k) for k in (0, 1)]
    ref = P.reference_partition()
    ref_loads = [sum(c for c, s in zip(costs, ref) if s == k) for k in (0, 1)]
    assert max(loads) <= max(ref_loads)  # never worse than the reference's encoder|decoder split
    four = P.partition(costs, 4)
    assert four == sorted(four) and set(four) == {0, 1, 2, 3}
    names = {f.name for f in P.flows()}
    assert {"skip0", "skip3", "main0", "main7"} <= names and len(names) == 12
    assert P.micro_batch_sizes(4, 2) == [2, 2] and P.micro_batch_sizes(5, 2) == [3, 2] and P.micro_batch_sizes(1, 2) == [1]


def _run(world, schedule, micro, port):
    env = dict(os.environ, DPT_SCHEDULE=schedule, DPT_MICRO=str(micro), DPT_BACKEND="gloo", OMP_NUM_THREADS="2")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", str(port), os.path.join(ROOT, "tests", "pipe_worker.py")]
    r = run_group(cmd, timeout=600, env=env, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    res = [json.loads(l.split("PIPERESULT ", 1)[1]) for l in r.stdout.splitlines() if "PIPERESULT " in l]
    assert len(res) == world
    return res


@pytest.mark.parametrize("world,schedule,micro,port", [(2, "gpipe", 2, 29701), (4, "gpipe", 4, 29702), (4, "1f1b", 4, 29703)])
def test_pipeline_equals_unpipelined(world, schedule, micro, port):
    for r in _run(world, schedule, micro, port):
        assert r["max_param_diff"] < 2e-4, r
        if r["losses"]:
            for a, b in zip(r["losses"], r["losses_ref"]):
                assert abs(a - b) < 1e-4, r
