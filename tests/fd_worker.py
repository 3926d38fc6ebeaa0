"""Worker for test_fd_exchange.py: every rank passes a pipe's read end to every peer (all-gather pattern), then rank 0
passes one more descriptor to everybody (broadcast pattern) - the two patterns MulticastBuffer uses for CUDA VMM and
multicast-object handles.  gloo, no GPU."""
import json
import os
import sys

import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.parallel.comm import _exchange_fds  # noqa: E402

dist.init_process_group("gloo")
rank, world = dist.get_rank(), dist.get_world_size()
peers = [r for r in range(world) if r != rank]
import tempfile  # noqa: E402

f = tempfile.TemporaryFile()  # unlinked file: only reachable through the descriptor that is passed around
f.write(f"hello from {rank}".encode())
f.flush()
got = _exchange_fds(rank, world, {p: f.fileno() for p in peers}, len(peers), None)
texts = {src: os.pread(fd, 64, 0).decode() for src, fd in got.items()}
ok_gather = sorted(texts) == peers and all(texts[p] == f"hello from {p}" for p in peers)
if rank == 0:
    f2 = tempfile.TemporaryFile()
    f2.write(b"multicast object")
    f2.flush()
    _exchange_fds(0, world, {p: f2.fileno() for p in peers}, 0, None)
    ok_bcast = True
else:
    got = _exchange_fds(rank, world, {}, 1, None)
    ok_bcast = list(got) == [0] and os.pread(got[0], 64, 0) == b"multicast object"
print("FDRESULT " + json.dumps({"rank": rank, "gather": ok_gather, "bcast": ok_bcast, "sources": sorted(texts)}), flush=True)
dist.destroy_process_group()
