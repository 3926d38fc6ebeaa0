#!/usr/bin/env python3
"""Train UNet on images and target masks - same command line as the reference's train.py:15-26.

    python3 train.py                                  single GPU
    python3 train.py -t DP                            single-process data parallel over all GPUs
    torchrun --standalone --nnodes=1 --nproc_per_node=N train.py -t DDP -b 2
    python3 train.py -t MP                            pipeline model parallel (spawns one worker per stage)

Additions (not in the reference): --variant / --classes / --bilinear select the north-star (milesial) model,
--stages / --micro-batches / --schedule configure the pipeline, --load finally does what its help text says,
--resume continues a run from its last epoch (weights + optimizer + scheduler + step; single GPU, DP and DDP).
Environment: DPT_IMAGE_SIZE=WxH and DPT_SYNTHETIC=N (synthetic dataset of N examples) for machines without Carvana.
"""
import argparse
import logging
import os
import warnings

import torch
import torch.distributed as dist

from distributedpytorch_b200.models import UNet
from distributedpytorch_b200.trainer import fit, fit_DDP, fit_DP, fit_MP
from distributedpytorch_b200.utils.checkpoint import load_checkpoint
from distributedpytorch_b200.utils.logging_utils import setup_logging
from distributedpytorch_b200.utils.loss import CEDiceLoss, Loss
from distributedpytorch_b200.utils.seed import set_seed

warnings.filterwarnings("ignore")

BACKEND = "nccl"
INIT_METHOD = "env://"


def get_args(argv=None):
    parser = argparse.ArgumentParser(description="Train UNet on images and target masks")
    parser.add_argument("--train-method", "-t", type=str, default="singleGPU", help="Training method: singleGPU | DP | DDP | MP")
    parser.add_argument("--validation", "-v", dest="val", type=float, default=10.0, help="Percentage of data used as validation")
    parser.add_argument("--load", "-l", type=str, default=False, help="Load model from a .pth file")
    parser.add_argument("--epochs", "-e", type=int, default=10, help="Number of epochs")
    parser.add_argument("--learning-rate", "--lr", type=float, default=1e-4, help="Learning rate", dest="lr")
    parser.add_argument("--batch-size", "-b", type=int, default=4, help="Batch size")
    parser.add_argument("--checkpoint", "-c", type=str, default=None, help="File name of the checkpoint to load")
    parser.add_argument("--seed", "-s", type=int, default=42, help="Set seed for reproducibility")
    parser.add_argument("--variant", type=str, default="ref", choices=["ref", "milesial"], help="UNet family member")
    parser.add_argument("--classes", type=int, default=None, help="Number of classes (milesial variant)")
    parser.add_argument("--bilinear", action="store_true", help="Bilinear up-sampling (milesial variant)")
    parser.add_argument("--stages", type=int, default=None, help="Pipeline stages for -t MP (default 2)")
    parser.add_argument("--micro-batches", type=int, default=2, help="Pipeline micro-batches for -t MP")
    parser.add_argument("--schedule", type=str, default="gpipe", choices=["gpipe", "1f1b"], help="Pipeline schedule")
    parser.add_argument("--resume", action="store_true",
                        help="Continue from checkpoints/<mode>.pth + <mode>.resume.pth (optimizer, scheduler, epoch, step)")
    parser.add_argument("--uint8-input", action="store_true",
                        help="Loaders hand over raw uint8 pixels/masks, normalised on the GPU (native engine; = DPT_UINT8_INPUT=1)")
    return parser.parse_args(argv)


def main(argv=None):
    args = get_args(argv)
    if args.uint8_input:
        os.environ["DPT_UINT8_INPUT"] = "1"
    set_seed(args.seed)
    rank = int(os.environ.get("RANK", "0"))
    setup_logging(args.train_method, rank=rank)
    logging.info("UNet for Carvana Image Masking (Segmentation)")
    model = UNet(pipe=args.train_method == "MP", variant=args.variant, n_classes=args.classes, bilinear=args.bilinear)
    if args.checkpoint is not None:
        load_checkpoint(model, args.checkpoint)
    elif args.load:
        load_checkpoint(model, args.load)
    criterion = Loss() if args.variant == "ref" else CEDiceLoss(model.n_classes)
    world_size = torch.cuda.device_count()
    common = dict(model=model, criterion=criterion, epochs=args.epochs, batch_size=args.batch_size, learning_rate=args.lr,
                  val_percent=args.val)
    method = args.train_method
    if os.environ.get("DPT_DEBUG_SKIP_OPS"):
        raise SystemExit("DPT_DEBUG_SKIP_OPS removes ops from the step program (scripts/ablate_step.py only); unset it to train")
    if method == "singleGPU":
        device = torch.device("cuda:0" if torch.cuda.is_available() else "cpu")
        logging.info(f"Using {device} for single-GPU training.")
        fit(resume=args.resume, **common)
    elif method == "DP":
        fit_DP(device_ids=list(range(world_size)) if world_size else None, resume=args.resume, **common)
    elif method == "DDP":
        backend = BACKEND if torch.cuda.is_available() else "gloo"
        kwargs = {"device_id": torch.device("cuda", int(os.environ.get("LOCAL_RANK", "0")))} if backend == "nccl" else {}
        if backend == "nccl":
            torch.cuda.set_device(int(os.environ.get("LOCAL_RANK", "0")))
        dist.init_process_group(backend=backend, init_method=INIT_METHOD, **kwargs)
        try:
            fit_DDP(world_size=dist.get_world_size(), resume=args.resume, **common)
        finally:
            dist.destroy_process_group()
    elif method == "MP":
        fit_MP(n_stages=args.stages, n_micro=args.micro_batches, schedule=args.schedule, **common)
    else:
        raise SystemExit(f"unknown --train-method {method!r} (expected singleGPU, DP, DDP or MP)")


if __name__ == "__main__":
    main()
