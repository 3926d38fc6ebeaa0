"""File-descriptor passing between ranks (AF_UNIX / SCM_RIGHTS): the host-side half of the VMM / NVLS multicast buffers
(distributedpytorch_b200/parallel/comm.py MulticastBuffer, csrc/comm/vmm.cu).  CPU only, gloo, 3 ranks."""
import json
import os
import sys

import torch

from proc_util import run_group

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_fd_exchange_three_ranks():
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=3", "--master-addr", "127.0.0.1",
           "--master-port", "29731", os.path.join(ROOT, "tests", "fd_worker.py")]
    r = run_group(cmd, timeout=240, env=dict(os.environ, OMP_NUM_THREADS="1"), cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-3000:]
    import re

    # the ranks share one stdout: two records may land on one line
    res = [json.loads(m) for m in re.findall(r"FDRESULT (\{.*?\})", r.stdout)]
    assert len(res) == 3
    for x in res:
        assert x["gather"] and x["bcast"], x
        assert x["sources"] == [p for p in range(3) if p != x["rank"]]


def test_multicast_query_without_gpu_is_false():
    from distributedpytorch_b200.parallel.comm import multicast_supported

    if not torch.cuda.is_available():
        assert multicast_supported(torch.device("cuda:0")) is False
