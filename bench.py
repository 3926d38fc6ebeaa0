#!/usr/bin/env python
"""Headline benchmark: UNet training images/sec (BASELINE.json metric) on N B200s of one node.

    python bench.py --gpus 1 --steps 20 --warmup 5
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W
    python bench.py --impl reference ...      # the unmodified reference (baseline/_ref) on its stock torch path

Config = the reference's default training configuration (train.py:20-22, utils/train_utils.py:26): reference
UNet (7.76 M params), batch 4 per GPU, 640x960 RGB images + binary masks, Adam(lr*world, wd 1e-8), loss =
BCE - log Dice, DDP for N > 1 (weak scaling: per-GPU batch fixed).  Synthetic data, random-init weights.

Timed region: exactly K steps (forward + loss + backward + gradient collective + optimizer) between
barrier + synchronize, CUDA events on the launching stream, max over ranks.  The per-step working set
(~1 GB of activations per image) is far larger than the 126 MB L2, so no explicit L2 flush is needed.
`e2e` repeats the measurement through the public step API with, every step, the host->device copy of that
step's batch from pinned memory and a device->host read of the loss.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import tempfile
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402

H, W = 640, 960
PER_GPU_BATCH = 4
METRIC = "UNet train images/sec (reference UNet 7.76M params, 640x960, batch 4/GPU, DDP weak scaling, max over ranks)"


def parse():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--impl", choices=["native", "reference"], default="native")
    ap.add_argument("--batch", type=int, default=PER_GPU_BATCH, help="per-GPU batch (default: the reference's 4)")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--model", choices=["ref", "milesial", "milesial-bilinear"], default="ref",
                    help="ref = the reference's UNet (headline); milesial* = the north-star BatchNorm UNet (31M / 17M params)")
    ap.add_argument("--amp", choices=["none", "bf16"], default="none",
                    help="with --impl reference: 'bf16' runs the reference's model under torch.autocast(bf16) + channels_last "
                         "(loss in fp32: nn.BCELoss refuses autocast) - the precision-matched comparator, BASELINE.md row R2; the "
                         "driver's reference arm is the stock fp32 path (default)")
    ap.add_argument("--torch-baseline", action="store_true",
                    help="with --model milesial*: run the same module on torch ops (bf16 autocast, channels_last) instead")
    return ap.parse_args()


class ClockSampler:
    """Samples nvidia-smi clocks / throttle reasons of one GPU every 200 ms while the timed region runs."""

    FIELDS = ("clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.hw_slowdown,"
              "clocks_event_reasons.hw_thermal_slowdown,clocks_event_reasons.sw_thermal_slowdown,"
              "clocks_event_reasons.sw_power_cap")

    def __init__(self, index: int):
        self.index = index
        self.proc = None
        self.path = None

    def start(self):
        try:
            fd, self.path = tempfile.mkstemp(suffix=".csv")
            os.close(fd)
            self.proc = subprocess.Popen(
                ["nvidia-smi", f"--id={self.index}", f"--query-gpu={self.FIELDS}", "--format=csv,noheader,nounits",
                 "-lms", "200"], stdout=open(self.path, "w"), stderr=subprocess.DEVNULL)
        except Exception:
            self.proc = None

    def stop(self) -> dict:
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.25)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=5)
        except Exception:
            self.proc.kill()
        sm, mx, reasons = [], [], set()
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        try:
            for line in open(self.path):
                parts = [p.strip() for p in line.split(",")]
                if len(parts) < 7:
                    continue
                try:
                    sm.append(float(parts[0]))
                    mx.append(float(parts[1]))
                except ValueError:
                    continue
                for n, v in zip(names, parts[3:7]):
                    if v.lower().startswith("active"):
                        reasons.add(n)
            os.unlink(self.path)
        except Exception:
            pass
        sm.sort()
        return {"sm_mhz": sm[len(sm) // 2] if sm else None, "sm_max_mhz": max(mx) if mx else None,
                "samples": len(sm), "reasons": sorted(reasons)}


def dist_setup(n_gpus: int):
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if n_gpus > 1 and world == 1:
        raise SystemExit("--gpus N > 1 must be launched with torch.distributed.run (one rank per GPU)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        dist.init_process_group("nccl", init_method="env://", device_id=torch.device("cuda", local))
    return rank, world, local


def barrier(world):
    if world > 1:
        dist.barrier()
    torch.cuda.synchronize()


def max_over_ranks(x: float, world: int) -> float:
    if world == 1:
        return x
    t = torch.tensor([x], dtype=torch.float64, device="cuda")
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def timed(step_fn, steps: int, world: int) -> float:
    """ms for `steps` calls of step_fn, device-timed, max over ranks."""
    barrier(world)
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        step_fn()
    e1.record()
    torch.cuda.synchronize()
    barrier(world)
    return max_over_ranks(e0.elapsed_time(e1), world)


def bench_config(per_gpu_batch: int, world: int) -> dict:
    """The benchmark configuration, identical (keys and values) for the native and the reference arm."""
    return {"model": "reference UNet (conv_block 32-64-128-256-512, 7,760,097 params)",
            "global_batch": per_gpu_batch * world, "per_gpu_batch": per_gpu_batch, "image": [3, H, W],
            "parallelism": "single" if world == 1 else f"ddp{world}",
            "optimizer": "Adam(lr=1e-4*world, weight_decay=1e-8)", "loss": "BCE - log Dice (utils/utils.py:9-25)",
            "l2": "per-step working set (>1 GB of activations) exceeds the 126 MB L2; no explicit flush"}


def synthetic_host_batches(batch: int, count: int, seed: int, uint8: bool = False):
    """`count` pinned host batches: fp32 images [B,3,H,W] in [0,1] and fp32 binary masks [B,H,W] (what the reference's
    loader yields), or - `uint8` - the same as raw 8-bit pixels and 8-bit masks, which the native engine normalises on
    the device."""
    g = torch.Generator().manual_seed(seed)
    out = []
    for _ in range(count):
        if uint8:
            img = torch.randint(0, 256, (batch, 3, H, W), generator=g, dtype=torch.uint8).pin_memory()
        else:
            img = torch.rand(batch, 3, H, W, generator=g).pin_memory()
        yy = torch.arange(H).view(1, H, 1).float()
        xx = torch.arange(W).view(1, 1, W).float()
        cy = (0.35 + 0.3 * torch.rand(batch, 1, 1, generator=g)) * H
        cx = (0.3 + 0.4 * torch.rand(batch, 1, 1, generator=g)) * W
        mask = (((yy - cy) / (0.2 * H)) ** 2 + ((xx - cx) / (0.25 * W)) ** 2) <= 1
        mask = (mask.to(torch.uint8) if uint8 else mask.float()).pin_memory()
        out.append((img, mask))
    return out


# ------------------------------------------------------------------------------------------------------
def run_native(args, rank, world, local):
    from distributedpytorch_b200.models import UNet
    from distributedpytorch_b200.ops import native

    dev = torch.device("cuda", local)
    torch.manual_seed(42)
    if args.model != "ref":
        return run_milesial(args, rank, world, local)
    model = UNet()
    lr = 1e-4 * world  # the reference scales lr by the world size in DDP (utils/train_utils.py:199)
    # raw 8-bit pixels and masks from the host, normalised by the input-packing kernel (DPT_BENCH_FLOAT_INPUT=1: fp32 as the
    # reference's loader yields them)
    in_dt = "float32" if os.environ.get("DPT_BENCH_FLOAT_INPUT", "0") == "1" else "uint8"
    if world == 1:
        from distributedpytorch_b200.engine import NativeUNetEngine

        stepper = NativeUNetEngine(model, args.batch, H, W, dev, lr=lr, input_dtype=in_dt, target_dtype=in_dt)
        parallelism = "single"
    else:
        from distributedpytorch_b200.parallel.ddp import FusedDDP

        stepper = FusedDDP(model, args.batch, H, W, dev, lr=lr, algo=os.environ.get("DPT_DDP_ALGO", "auto"),
                           overlap=bool(int(os.environ.get("DPT_DDP_OVERLAP", "0"))), input_dtype=in_dt, target_dtype=in_dt)
        parallelism = f"ddp{world}"
    host = synthetic_host_batches(args.batch, 2, seed=1234 + rank, uint8=in_dt == "uint8")
    stepper.load_batch(*host[0])
    torch.cuda.synchronize()

    # ---- device-timed steps on pre-staged inputs ----
    for _ in range(args.warmup):
        stepper.train_step()
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    before = native.launch_count()
    ms = timed(lambda: stepper.train_step(), args.steps, world)
    launches = native.launch_count() - before
    clocks = sampler.stop() if rank == 0 else {}
    loss = float(stepper.loss_state[4].item())

    # ---- end to end: pinned host batch -> device every step, loss read back every step ----
    e2e = None
    if not args.no_e2e:
        k = [0]

        def e2e_step():
            # batch k was handed to the copy stream during step k-1 (pinned host -> device, overlapped with compute);
            # this call consumes it, launches step k, starts the copy of batch k+1 and reads back the loss of step k.
            loss_t = stepper.train_step()
            k[0] += 1
            stepper.prefetch(*host[k[0] & 1])
            return float(loss_t.item())  # device -> host read of the step's result (4 bytes), syncs like the reference

        stepper.prefetch(*host[0])
        for _ in range(max(2, args.warmup // 2)):
            e2e_step()
        ms_e2e = timed(e2e_step, args.steps, world)
        h2d = sum(t.numel() * t.element_size() for t in host[0])  # per GPU
        e2e = {"value": round(args.batch * world * args.steps / (ms_e2e / 1e3), 2), "unit": "images/s",
               "ms_per_step": round(ms_e2e / args.steps, 3), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4,
               "input": f"{in_dt} images + {in_dt} masks from pinned host memory, copied on a side stream straight into the input "
                        "buffers as soon as the step in flight has read them (program events), no staging hop"}
    if rank == 0:
        value = args.batch * world * args.steps / (ms / 1e3)
        out = {
            "metric": METRIC, "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": "native",
            "config": bench_config(args.batch, world),
            "notes": "fp32 master weights, bf16 activations / kernel-side weights; fused all-reduce + Adam for N > 1",
            "clocks": clocks, "e2e": e2e, "gpu_launches": int(launches), "final_loss": round(loss, 5),
        }
        emit(out)


# ------------------------------------------------------------------------------------------------------
def run_milesial(args, rank, world, local):
    """Secondary benchmark (BASELINE.json config 5): the north-star BatchNorm UNet, native engine or the same module on
    torch ops with bf16 autocast + channels_last (the reference has no code for this model)."""
    from distributedpytorch_b200.models import UNet
    from distributedpytorch_b200.ops import native
    from distributedpytorch_b200.utils.loss import CEDiceLoss

    dev = torch.device("cuda", local)
    bilinear = args.model == "milesial-bilinear"
    model = UNet(variant="milesial", n_classes=2, bilinear=bilinear)
    host = synthetic_host_batches(args.batch, 2, seed=1234 + rank)
    images, masks = host[0][0].to(dev), host[0][1].to(dev)
    if args.torch_baseline:
        model = model.to(dev).to(memory_format=torch.channels_last)
        if world > 1:
            model = torch.nn.parallel.DistributedDataParallel(model, device_ids=[local])
        opt = torch.optim.RMSprop(model.parameters(), lr=1e-5, weight_decay=1e-8, momentum=0.999, foreach=True)
        crit = CEDiceLoss(2)
        x = images.contiguous(memory_format=torch.channels_last)
        t = masks.long()

        def step():
            with torch.autocast("cuda", dtype=torch.bfloat16):
                loss = crit(model(x), t)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()

        impl, before = "torch bf16 autocast + channels_last (same module)", 0
    else:
        if world == 1:
            from distributedpytorch_b200.engine.milesial import NativeMilesialEngine

            stepper = NativeMilesialEngine(model, args.batch, H, W, dev, lr=1e-5)
        else:
            from distributedpytorch_b200.parallel.ddp import FusedDDP

            stepper = FusedDDP(model, args.batch, H, W, dev, lr=1e-5)
        stepper.load_batch(images, masks.to(torch.uint8))
        step = stepper.train_step
        impl, before = "native", native.launch_count()
    for _ in range(args.warmup):
        step()
    if not args.torch_baseline:
        before = native.launch_count()
    ms = timed(step, args.steps, world)
    launches = 0 if args.torch_baseline else native.launch_count() - before
    if rank == 0:
        emit({"metric": "UNet (north-star BatchNorm variant) train images/sec", "value": round(args.batch * world * args.steps / (ms / 1e3), 2),
              "unit": "images/s", "n_gpus": world, "steps": args.steps, "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3),
              "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "bf16", "data": "synthetic", "impl": impl,
              "config": {"model": args.model, "params": sum(p.numel() for p in model.parameters()), "per_gpu_batch": args.batch,
                         "global_batch": args.batch * world, "image": [3, H, W], "optimizer": "RMSprop(momentum 0.999, wd 1e-8)",
                         "loss": "cross-entropy + Dice"},
              "gpu_launches": int(launches)})


# ------------------------------------------------------------------------------------------------------
def run_reference(args, rank, world, local):
    ref_dir = os.path.join(ROOT, "baseline", "_ref")
    if not os.path.isfile(os.path.join(ref_dir, "model", "unet_model.py")):
        if rank == 0:
            emit({"impl": "reference", "unavailable": "baseline/_ref is missing (run baseline/install_reference.sh); "
                  "the reference has no setup.py/pyproject so pip cannot install it"})
        return
    import types

    if "matplotlib" not in sys.modules:  # imported by the reference's utils/utils.py:1, not installed on this image
        mpl = types.ModuleType("matplotlib")
        mpl.pyplot = types.ModuleType("matplotlib.pyplot")
        sys.modules["matplotlib"] = mpl
        sys.modules["matplotlib.pyplot"] = mpl.pyplot
    sys.path.insert(0, ref_dir)
    saved = {k: sys.modules.pop(k) for k in list(sys.modules) if k == "utils" or k.startswith("utils.") or k == "model"
             or k.startswith("model.")}
    from model.unet_model import UNet as RefUNet  # the reference's own classes, unmodified (model/__init__.py is empty)
    from utils.utils import Loss as RefLoss, set_seed as ref_set_seed
    sys.modules.update(saved)

    dev = torch.device("cuda", local)
    ref_set_seed(42)
    model = RefUNet()
    criterion = RefLoss()
    batch = args.batch
    amp = args.amp == "bf16"
    if amp:
        model = model.to(memory_format=torch.channels_last)
    if world == 1:
        model = model.cuda(local)  # train.py:49
        lr = 1e-4
    else:
        from torch.nn.parallel import DistributedDataParallel as DDP  # utils/train_utils.py:195-199

        model.to(local)
        model = DDP(model, device_ids=[local], output_device=local)
        lr = 1e-4 * world
    optimizer = torch.optim.Adam(model.parameters(), lr=lr, weight_decay=1e-8)
    host = synthetic_host_batches(batch, 2, seed=1234 + rank)
    dev_batches = [(i.to(dev), m.to(dev)) for i, m in host]

    def step(images, masks):
        # the body of the reference's training loop, utils/train_utils.py:60-70 / :213-225
        images = images.to(dev, non_blocking=True).to(torch.float32)
        true_masks = masks.to(dev, non_blocking=True).to(torch.float32).unsqueeze(1)
        if amp:
            with torch.autocast("cuda", dtype=torch.bfloat16):
                pred = model(images.contiguous(memory_format=torch.channels_last))
            pred = pred.float()
        else:
            pred = model(images)
        loss = criterion(pred, true_masks)
        optimizer.zero_grad()
        val = loss.item()
        (batch * loss).backward()
        optimizer.step()
        return val

    for _ in range(args.warmup):
        step(*dev_batches[0])
    sampler = ClockSampler(local)
    if rank == 0:
        sampler.start()
    ms = timed(lambda: step(*dev_batches[0]), args.steps, world)
    clocks = sampler.stop() if rank == 0 else {}
    k = [0]

    def e2e_step():
        img, mask = host[k[0] & 1]
        k[0] += 1
        return step(img, mask)

    e2e = None
    if not args.no_e2e:
        for _ in range(2):
            e2e_step()
        ms_e2e = timed(e2e_step, args.steps, world)
        h2d = host[0][0].numel() * 4 + host[0][1].numel() * 4
        e2e = {"value": round(batch * world * args.steps / (ms_e2e / 1e3), 2), "unit": "images/s",
               "ms_per_step": round(ms_e2e / args.steps, 3), "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": 4}
    if rank == 0:
        value = batch * world * args.steps / (ms / 1e3)
        emit({
            "metric": METRIC, "value": round(value, 2), "unit": "images/s", "n_gpus": world, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": round(ms / args.steps, 3), "higher_is_better": True,
            "scaling": "weak", "vs_baseline": None,
            "dtype": "bf16 autocast (fp32 loss and optimizer)" if amp else "fp32 (torch default: TF32 allowed in cuDNN convolutions)",
            "data": "synthetic", "impl": "reference+bf16-autocast" if amp else "reference",
            "config": bench_config(batch, world),
            "notes": "unmodified model/ + utils/utils.py from baseline/_ref; step body of utils/train_utils.py:60-70 incl. "
                     "loss.item() every step; torch DDP (NCCL) for N > 1" + ("; model under bf16 autocast + channels_last" if amp else ""),
            "clocks": clocks, "e2e": e2e, "gpu_launches": 0,
        })


def emit(obj) -> None:
    """Write the ONE result line to the real stdout (everything else, e.g. NCCL's version banner, goes to stderr)."""
    os.write(_REAL_STDOUT, (json.dumps(obj) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    args = parse()
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)  # libraries that print to fd 1 (NCCL banner) must not pollute the JSON line
    if not torch.cuda.is_available():
        emit({"impl": args.impl, "unavailable": "no CUDA device"})
        return
    if os.environ.get("DPT_DEBUG_SKIP_OPS"):
        # the ablation switch of scripts/ablate_step.py removes ops from the step program: never a benchmark number
        raise SystemExit("bench.py refuses to run with DPT_DEBUG_SKIP_OPS set (work would be skipped inside the timed region)")
    rank, world, local = dist_setup(args.gpus)
    try:
        if args.impl == "reference":
            run_reference(args, rank, world, local)
        else:
            run_native(args, rank, world, local)
    finally:
        if world > 1 and dist.is_initialized():
            dist.destroy_process_group()


if __name__ == "__main__":
    main()
