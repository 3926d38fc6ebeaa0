"""Device time and roofline fraction of the fused gradient collective (reduce-scatter + scale + Adam + all-gather in one
kernel) on N GPUs, one rank per GPU:

    python -m torch.distributed.run --nproc-per-node N scripts/bench_collective.py [--mb 31]

Roofline (B200_PROFILING.md): the slower of  bytes over NVLink / 770 GB/s per direction per GPU  and  optimizer bytes over
measured HBM bandwidth.  Two-shot moves (N-1)/N * S in and (N-1)/N * S out per GPU; one-shot moves (N-1) * S in.
The optimizer touches p, g, m, v (read) and p, m, v (write) for the rank's slice (two-shot) or the whole arena (one-shot).
Also times the same job done the reference's way: NCCL all-reduce (torch.distributed) + torch's fused multi-tensor Adam.
"""
import argparse
import ctypes as C
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.ops import native as N  # noqa: E402
from distributedpytorch_b200.parallel.comm import ProcessTeam, multicast_supported  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--mb", type=float, default=31.04, help="gradient arena size in MB (fp32); default = the reference UNet")
ap.add_argument("--iters", type=int, default=20)
ap.add_argument("--stress", type=int, default=0,
                help="instead of timing: run this many back-to-back fused collectives with one rank delayed by a random spin "
                     "before some of them (exercises the epoch-flag barrier, which compute-sanitizer cannot check here) and "
                     "verify after every 100 that all ranks hold identical weights equal to a torch oracle")
args = ap.parse_args()

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
dev = torch.device("cuda", local)
n = int(args.mb * 1e6 / 4) // 4 * 4
team = ProcessTeam(dev)
use_mc = world > 1 and multicast_supported(dev)  # NVLS: arenas on VMM allocations bound to a multicast object
sp, sg = (team.mc_buffer(n * 4), team.mc_buffer(n * 4)) if use_mc else (team.buffer(n * 4), team.buffer(n * 4))
p, g = sp.tensor(torch.float32, n), sg.tensor(torch.float32, n)
p.normal_()
g.normal_()
m, v = torch.zeros(n, device=dev), torch.zeros(n, device=dev)
state = torch.zeros(8, device=dev)
L = N.lib()
N.check(L.dpt_opt_init(state.data_ptr(), 1e-4, 1, 0.9, 0.999, N.stream_ptr()))
peaks = {"hbm_gbs": 6540.8}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
except Exception:
    pass
LINK = 770e9


def timed(fn):
    for _ in range(3):
        fn()
    dist.barrier()
    torch.cuda.synchronize()
    ts = []
    for _ in range(args.iters):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    t = torch.tensor([ts[len(ts) // 2]], device=dev, dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


if args.stress:
    # ---- stress of the cross-GPU barrier protocol ---------------------------------------------------------------
    import random

    rng = random.Random(1234)  # same sequence on every rank
    results = []
    algos_s = [("two_shot", N.AR_TWO_SHOT), ("one_shot", N.AR_ONE_SHOT)] + ([("multimem", N.AR_MULTIMEM)] if use_mc else [])
    for algo_name, algo in algos_s:
        torch.manual_seed(5)
        p.copy_(torch.randn(n, device=dev))  # same start everywhere
        m.zero_(), v.zero_()
        N.check(L.dpt_opt_init(state.data_ptr(), 1e-3, 0, 0.9, 0.999, N.stream_ptr()))
        d = N.FusedArDesc()
        for r in range(world):
            d.grads[r], d.params[r] = sg.ptrs[r], sp.ptrs[r]
        if use_mc:
            d.mc_grads, d.mc_params = sg.mc_ptr, sp.mc_ptr
        d.state0, d.state1, d.opt_state = m.data_ptr(), v.data_ptr(), state.data_ptr()
        d.n, d.algo, d.optimizer, d.blocks = n, algo, N.OPT_ADAM, 64
        d.eps, d.weight_decay, d.grad_scale = 1e-8, 1e-8, 1.0 / world
        ok = True
        for it in range(args.stress):
            gen = torch.Generator(device=dev).manual_seed(1000 * it + rank)
            g.copy_(torch.randn(n, device=dev, generator=gen))
            check = (it + 1) % 100 == 0
            slow, spin = rng.randrange(world), rng.choice([0, 0, 0, 20000, 200000, 2000000])
            if rank == slow and spin:
                torch.cuda._sleep(spin)  # this rank's stream reaches the collective late
            N.check(L.dpt_opt_tick(state.data_ptr(), N.stream_ptr()))
            N.check(L.dpt_fused_allreduce_opt(C.byref(team.struct), C.byref(d), N.stream_ptr()))
            if check:
                # a rank that read a peer's gradients too early (or whose weights were overwritten too early) ends up
                # with different weights than its peers: compare the replicas bit for bit
                torch.cuda.synchronize()
                flat = [torch.empty_like(p) for _ in range(world)]
                dist.all_gather(flat, p)
                ok = ok and all(torch.equal(flat[0], f) for f in flat) and bool(torch.isfinite(p).all())
        torch.cuda.synchronize()
        results.append({"algo": algo_name, "steps": args.stress, "world": world, "replicas_identical_every_100": ok,
                        "arena_mb": n * 4 / 1e6})
    if rank == 0:
        for r in results:
            print("STRESS " + json.dumps(r), flush=True)
    dist.barrier()
    dist.destroy_process_group()
    sys.exit(0 if all(r["replicas_identical_every_100"] for r in results) else 1)

rows = []
algos = [("two_shot", N.AR_TWO_SHOT), ("one_shot", N.AR_ONE_SHOT)] + ([("multimem", N.AR_MULTIMEM)] if use_mc else [])
for algo_name, algo in algos:
    for blocks in (32, 64, 128):
        d = N.FusedArDesc()
        for r in range(world):
            d.grads[r], d.params[r] = sg.ptrs[r], sp.ptrs[r]
        if use_mc:
            d.mc_grads, d.mc_params = sg.mc_ptr, sp.mc_ptr
        d.state0, d.state1, d.opt_state = m.data_ptr(), v.data_ptr(), state.data_ptr()
        d.n, d.algo, d.optimizer, d.blocks = n, algo, N.OPT_ADAM, blocks
        d.eps, d.weight_decay, d.grad_scale = 1e-8, 1e-8, 1.0 / world
        ms = timed(lambda: N.check(L.dpt_fused_allreduce_opt(C.byref(team.struct), C.byref(d), N.stream_ptr())))
        S = n * 4
        if algo != N.AR_ONE_SHOT:
            link_bytes = S * (world - 1) / world  # per direction per GPU
            opt_bytes = 7 * S / world + S / world * 0  # p,g,m,v read + p,m,v written for the slice
        else:
            link_bytes = S * (world - 1)
            opt_bytes = 7 * S
        t_roof = max(link_bytes / LINK, opt_bytes / (peaks["hbm_gbs"] * 1e9)) * 1e3
        rows.append({"kernel": f"fused allreduce+adam {algo_name}", "world": world, "blocks": blocks, "arena_mb": S / 1e6,
                     "ms": round(ms, 4), "roof_ms": round(t_roof, 4), "frac_of_roof": round(t_roof / ms, 3),
                     "link_gbs_per_dir": round(link_bytes / ms / 1e6, 1)})

# the reference's way: NCCL all-reduce + torch fused Adam on the same arena
pr = torch.nn.Parameter(p.clone())
opt = torch.optim.Adam([pr], lr=1e-4, weight_decay=1e-8, fused=True)
gr = g.clone()


def ref_step():
    dist.all_reduce(gr)
    gr.div_(world)
    pr.grad = gr
    opt.step()


ms = timed(ref_step)
rows.append({"kernel": "NCCL all_reduce + div + torch fused Adam (reference path)", "world": world, "arena_mb": n * 4 / 1e6,
             "ms": round(ms, 4)})
if rank == 0:
    print(json.dumps({"nvls_multicast": bool(use_mc), "world": world}), flush=True)
    for r in rows:
        print(json.dumps(r), flush=True)
dist.barrier()
dist.destroy_process_group()
