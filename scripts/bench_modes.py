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
ap.add_argument("--variant", default="ref", choices=["ref", "milesial", "milesial-bilinear"])
ap.add_argument("--reserve", type=int, default=-1, help="mp: SMs left to NCCL (default: DPT_SM_RESERVE or 16)")
ap.add_argument("--nccl-channels", type=int, default=0, help="mp: NCCL_MAX_NCHANNELS / NCCL_MAX_P2P_NCHANNELS (default: reserve / 2)")
ap.add_argument("--flops-cut", action="store_true", help="mp: partition by forward GFLOP instead of measured block times")
ap.add_argument("--algo", default="auto", help="dp: collective algorithm of the fused reduce+update kernel")
ap.add_argument("--comm-blocks", type=int, default=64)
args = ap.parse_args()


def make_model():
    from distributedpytorch_b200.models import UNet

    if args.variant == "ref":
        return UNet()
    return UNet(variant="milesial", n_classes=2, bilinear=args.variant.endswith("bilinear"))


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
        model = make_model()
        dp = NativeDP(model, B, H, W, list(range(n)), lr=1e-4 if args.variant == "ref" else 1e-5,
                      optimizer="adam" if args.variant == "ref" else "rmsprop", algo=args.algo, comm_blocks=args.comm_blocks)
        x = torch.rand(B, 3, H, W, device="cuda:0")
        t = (torch.rand(B, H, W, device="cuda:0") > 0.5)
        dp.load_batch(x, t.float() if args.variant == "ref" else t.to(torch.uint8))
        for _ in range(args.warmup):
            dp.train_step()
        ms = timed(dp.train_step, args.steps, dp.synchronize)
        # what the fused collectives must move per step and GPU over NVLink, against the 770 GB/s measured per direction:
        # gradients in (reduce-scatter half) + weights out (all-gather half) = 2 * (N-1)/N * arena; input chunk; outputs
        n_g = len(dp.devices)
        arena = dp.arena_numel * 4
        wire = 2 * (n_g - 1) / n_g * arena
        print(json.dumps({"mode": f"native DP ({args.variant})", "gpus": n_g, "global_batch": B, "ms_per_step": ms / args.steps,
                          "images_per_s": B * args.steps / (ms / 1e3), "algo": dp.algo, "arena_bytes": arena,
                          "nvlink_bytes_per_gpu_per_dir": wire, "nvlink_floor_ms": wire / 770e9 * 1e3}))
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
        from distributedpytorch_b200.trainer.fit_mp import configure_pipeline_runtime

        if args.reserve >= 0:
            os.environ["DPT_SM_RESERVE"] = str(args.reserve)
        if args.nccl_channels > 0:
            os.environ["NCCL_MAX_NCHANNELS"] = os.environ["NCCL_MAX_P2P_NCHANNELS"] = str(args.nccl_channels)
        configure_pipeline_runtime()
        B = args.batch or 4 * args.micro // 2
        model = make_model()
        costs = model.stage_functions() if args.flops_cut else model.stage_costs()
        stage_of = P.partition([c for _, c in costs], world)
        dev = torch.device("cuda", local)
        groups = P.make_pair_groups(stage_of, list(range(world)), device=dev)
        lo, hi = stage_of.index(rank), len(stage_of) - stage_of[::-1].index(rank)
        sizes = P.micro_batch_sizes(B, args.micro)
        be = NativeStage(model, lo, hi, sizes, H, W, dev, lr=1e-4, loss_scale=float(B), whole_batch=args.schedule == "gpipe")
        stage = P.PipelineStage(be, rank, stage_of, len(sizes), args.schedule, groups)
        off = 0
        for mb, b in enumerate(sizes):
            m = torch.rand(b, H, W) > 0.5
            be.load_micro_batch(mb, torch.rand(b, 3, H, W) if rank == 0 else None,
                                (m.float() if args.variant == "ref" else m.to(torch.uint8)) if rank == world - 1 else None)
            off += b

        def sync():
            torch.cuda.synchronize()
            dist.barrier()

        for _ in range(args.warmup):
            stage.train_step()
        ms = timed(stage.train_step, args.steps, sync)
        t = torch.tensor([ms], device=dev, dtype=torch.float64)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        # second pass with CUDA events around every compute segment: busy ms per stage; step - busy = bubble + exposed transfer
        stage.timeline = True
        ms_tl = timed(stage.train_step, args.steps, sync)
        tl = stage.timeline_summary(args.steps)
        busy = torch.tensor([tl["busy_ms"], tl["fwd_ms"], tl["bwd_ms"]], device=dev, dtype=torch.float64)
        allbusy = [torch.zeros_like(busy) for _ in range(world)]
        dist.all_gather(allbusy, busy)
        if rank == 0:
            step = t.item() / args.steps
            per_stage = [round(float(b[0]), 3) for b in allbusy]
            print(json.dumps({"mode": f"native pipeline ({args.variant}) {world} stages {stage_of}, {len(sizes)} micro-batches, "
                                      f"{args.schedule}", "gpus": world, "global_batch": B, "ms_per_step": step,
                              "images_per_s": B * args.steps / (t.item() / 1e3),
                              "sm_reserve": int(os.environ.get("DPT_SM_RESERVE", "16")),
                              "nccl_channels": int(os.environ.get("NCCL_MAX_NCHANNELS", "0")),
                              "busy_ms_per_stage": per_stage, "fwd_ms_per_stage": [round(float(b[1]), 3) for b in allbusy],
                              "bwd_ms_per_stage": [round(float(b[2]), 3) for b in allbusy],
                              "ideal_ms": round(max(per_stage) * (len(sizes) + world - 1) / len(sizes), 3),
                              "idle_ms_critical_stage": round(ms_tl / args.steps - max(per_stage), 3)}))
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
