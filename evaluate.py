#!/usr/bin/env python3
"""Validation.  `evaluate(model, dataloader, criterion) -> float` is the reference's library function
(evaluate.py:6-25); run as a script it evaluates a checkpoint on the validation split and prints the mean loss and
the Dice score (the reference's evaluate.py has no command line of its own).

    python3 evaluate.py -c singleGPU [-b 4] [-v 10]
"""
import argparse

import torch
from torch.utils.data import DataLoader

from distributedpytorch_b200.models import UNet
from distributedpytorch_b200.trainer.evaluate import evaluate as _evaluate, evaluate_dice
from distributedpytorch_b200.trainer.fit import _dataset, _size
from distributedpytorch_b200.trainer.steppers import make_stepper
from distributedpytorch_b200.utils import data as D
from distributedpytorch_b200.utils.checkpoint import load_checkpoint
from distributedpytorch_b200.utils.loss import CEDiceLoss, Loss


def evaluate(model, dataloader, criterion) -> float:
    """Mean of `criterion` over `dataloader` without gradients; leaves the model in train mode."""
    return _evaluate(model, dataloader, criterion)


def main(argv=None):
    ap = argparse.ArgumentParser(description="Evaluate a UNet checkpoint on the validation split")
    ap.add_argument("--checkpoint", "-c", type=str, required=True)
    ap.add_argument("--batch-size", "-b", type=int, default=4)
    ap.add_argument("--validation", "-v", dest="val", type=float, default=10.0)
    ap.add_argument("--variant", type=str, default="ref", choices=["ref", "milesial"])
    ap.add_argument("--classes", type=int, default=None)
    ap.add_argument("--bilinear", action="store_true")
    args = ap.parse_args(argv)
    model = UNet(variant=args.variant, n_classes=args.classes, bilinear=args.bilinear)
    load_checkpoint(model, args.checkpoint)
    newsize = _size()
    _, val_set = D.split_dataset(_dataset(newsize), args.val, seed=0)
    loader = DataLoader(val_set, batch_size=args.batch_size, shuffle=False, drop_last=True)
    device = torch.device("cuda:0" if torch.cuda.is_available() else "cpu")
    criterion = Loss() if args.variant == "ref" else CEDiceLoss(model.n_classes)
    stepper, _ = make_stepper(model, criterion, args.batch_size, newsize, device, lr=0.0)
    loss = _evaluate(stepper, loader)
    dice = evaluate_dice(stepper, loader, n_classes=model.n_classes)
    print(f"validation loss {loss:.6f}  dice {dice:.4f}  ({len(val_set)} examples)")
    return loss, dice


if __name__ == "__main__":
    main()
