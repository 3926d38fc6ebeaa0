"""Validation loop (reference: evaluate.py:6-25): mean of the criterion over the validation loader, computed
without gradients, model left in train mode afterwards.  One host sync at the end instead of `loss.item()` per
batch.  `evaluate_dice` adds the Dice-score evaluation the north star asks for (not in the reference)."""
from __future__ import annotations

from typing import Iterable

import torch

from ..utils.loss import dice_score


def evaluate(stepper, dataloader: Iterable, criterion=None) -> float:
    """Mean validation loss.  `stepper` is any object with evaluate_batch(images, masks) (native engine, DDP/DP/
    pipeline stepper, TorchStepper) or a bare nn.Module together with `criterion`."""
    module = getattr(stepper, "model", None) if not isinstance(stepper, torch.nn.Module) else stepper
    if module is not None:
        module.eval()
    total, count = None, 0
    for batch in dataloader:
        images, masks = batch["image"], batch["mask"]
        if isinstance(stepper, torch.nn.Module):
            dev = next(stepper.parameters()).device
            with torch.no_grad():
                out = stepper(images.to(dev).float())
                loss = criterion(out, masks.to(dev).float().reshape(out.shape))
        else:
            if getattr(stepper, "B", None) is not None and images.shape[0] != stepper.B:
                continue  # fixed-shape native engine: the loader uses drop_last=True like the reference's val loader
            loss = stepper.evaluate_batch(images, masks)
        loss = loss.detach().float().reshape(())
        total = loss.clone() if total is None else total + loss
        count += 1
    if module is not None:
        module.train()
    if count == 0:
        return float("nan")
    return float((total / count).item())


@torch.no_grad()
def evaluate_dice(stepper, dataloader: Iterable, n_classes: int = 1) -> float:
    """Mean Dice score of the hard predictions."""
    scores = []
    for batch in dataloader:
        images, masks = batch["image"], batch["mask"]
        if getattr(stepper, "B", None) is not None and images.shape[0] != stepper.B:
            continue
        out = stepper.forward(images) if not isinstance(stepper, torch.nn.Module) else stepper(images)
        scores.append(dice_score(out.float().cpu(), masks.cpu(), n_classes))
    return float(torch.stack(scores).mean().item()) if scores else float("nan")
