"""Throughput of the single-process modes on all visible GPUs: native DataParallel vs
torch.nn.DataParallel on the reference model, native pipeline (GPipe / 1F1B, 2..N stages) vs the reference's
UNet(pipe=True).  Device-timed, JSON lines on stdout.

    python scripts/bench_modes.py --mode dp  [--batch-per-gpu 4]
    python scripts/bench_modes.py --mode dp-ref
    python scripts/bench_modes.py --mode mp-ref
    python -m torch.distributed.run --nproc-per-node S scripts/bench_modes.py --mode mp --micro 4 --schedule gpipe
"""
import argparse
import json
import os
import sys
import types

import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
H, W = 640, 960

ap = argparse.ArgumentParser()
ap.add_argument("--mode", required=True, choices=["dp", "dp-ref", "mp", "mp-ref"])
ap.add_argument("--batch-per-gpu", type=int, default=4)
ap.add_argument("--batch", type=int, default=0, help="global batch (default: batch-per-gpu x GPUs; mp: 4 x micro)")
ap.add_argument("--steps", type=int, default=10)
ap.add_argument("--warmup", type=int, default=3)
ap.add_argument("--micro", type=int, default=2)
ap.add_argument("--schedule", default="gpipe")
args = ap.parse_args()


def timed(fn, steps, sync):
    sync()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        fn()
    e1.record()
    sync()
    return e0.elapsed_time(e1)


def load_reference():
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if "matplotlib" not in sys.modules:
        mpl = types.ModuleType("matplotlib")
        mpl.pyplot = types.ModuleType("matplotlib.pyplot")
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = mpl.pyplot
    sys.path.insert(0, ref_dir)
    from model.unet_model import UNet as RefUNet
    from utils.utils import Loss as RefLoss
    return RefUNet, RefLoss


def main():
    n = torch.cuda.device_count()
    if args.mode == "dp":
        from distributedpytorch_b200.models import UNet
        from distributedpytorch_b200.parallel.dp import NativeDP

        B = args.batch or args.batch_per_gpu * n
        dp = NativeDP(UNet(), B, H, W, list(range(n)), lr=1e-4)
        x = torch.rand(B, 3, H, W, device="cuda:0")
        t = (torch.rand(B, H, W, device="cuda:0") > 0.5).float()
        dp.load_batch(x, t)
        for _ in range(args.warmup):
            dp.train_step()
        ms = timed(dp.train_step, args.steps, dp.synchronize)
        print(json.dumps({"mode": "native DP", "gpus": len(dp.devices), "global_batch": B, "ms_per_step": ms / args.steps,
                          "images_per_s": B * args.steps / (ms / 1e3)}))
    elif args.mode == "dp-ref":
        RefUNet, RefLoss = load_reference()
        B = args.batch or args.batch_per_gpu * n
        model = torch.nn.DataParallel(RefUNet().cuda(0), device_ids=list(range(n)))  # with the reference's two DP bugs fixed
        crit = RefLoss()
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, weight_decay=1e-8)
        x = torch.rand(B, 3, H, W, device="cuda:0")
        t = (torch.rand(B, 1, H, W, device="cuda:0") > 0.5).float()

        def step():
            loss = crit(model(x), t)
            opt.zero_grad()
            loss.item()
            (B * loss).backward()
            opt.step()

        for _ in range(args.warmup):
            step()
        ms = timed(step, args.steps, torch.cuda.synchronize)
        print(json.dumps({"mode": "torch DataParallel (reference model)", "gpus": n, "global_batch": B,
                          "ms_per_step": ms / args.steps, "images_per_s": B * args.steps / (ms / 1e3)}))
    elif args.mode == "mp-ref":
        RefUNet, RefLoss = load_reference()
        B = args.batch or 4
        import contextlib
        import io

        with contextlib.redirect_stdout(io.StringIO()):
            model = RefUNet(pipe=True)  # prints the decoder (model/unet_model.py:21)
        crit = RefLoss()
        opt = torch.optim.Adam(model.parameters(), lr=1e-4, weight_decay=1e-8)
        x = torch.rand(B, 3, H, W, device="cuda:0")
        t = (torch.rand(B, 1, H, W, device="cuda:0") > 0.5).float()

        def step():
            loss = crit(model(x), t)
            opt.zero_grad()
            loss.item()
            (B * loss).backward()
            opt.step()

        for _ in range(args.warmup):
            step()
        ms = timed(step, args.steps, lambda: (torch.cuda.synchronize(0), torch.cuda.synchronize(1)))
        print(json.dumps({"mode": "reference UNet(pipe=True), 2 stages x 2 micro-batches", "gpus": 2, "global_batch": B,
                          "ms_per_step": ms / args.steps, "images_per_s": B * args.steps / (ms / 1e3)}))
    else:  # native pipeline, one rank per stage
        import torch.distributed as dist

        from distributedpytorch_b200.engine.stage_engine import NativeStage
        from distributedpytorch_b200.models import UNet
        from distributedpytorch_b200.parallel import pipe as P

        rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local)
        dist.init_process_group("nccl")
        B = args.batch or 4 * args.micro // 2
        model = UNet()
        stage_of = P.partition([c for _, c in model.stage_functions()], world)
        dev = torch.device("cuda", local)
        groups = P.make_pair_groups(stage_of, list(range(world)), device=dev)
        lo, hi = stage_of.index(rank), len(stage_of) - stage_of[::-1].index(rank)
        sizes = P.micro_batch_sizes(B, args.micro)
        be = NativeStage(model, lo, hi, sizes, H, W, dev, lr=1e-4, loss_scale=float(B), whole_batch=args.schedule == "gpipe")
        stage = P.PipelineStage(be, rank, stage_of, len(sizes), args.schedule, groups)
        off = 0
        for mb, b in enumerate(sizes):
            be.load_micro_batch(mb, torch.rand(b, 3, H, W) if rank == 0 else None,
                                (torch.rand(b, H, W) > 0.5).float() if rank == world - 1 else None)
            off += b

        def sync():
            torch.cuda.synchronize()
            dist.barrier()

        for _ in range(args.warmup):
            stage.train_step()
        ms = timed(stage.train_step, args.steps, sync)
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        if rank == 0:
            print(json.dumps({"mode": f"native pipeline {world} stages {stage_of}, {len(sizes)} micro-batches, {args.schedule}",
                              "gpus": world, "global_batch": B, "ms_per_step": t.item() / args.steps,
                              "images_per_s": B * args.steps / (t.item() / 1e3)}))
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
