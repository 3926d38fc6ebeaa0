"""Losses and evaluation metrics.

`Loss` is the reference criterion (utils/utils.py:9-25): ``BCELoss(p, t) - dice_weight * log(2 I / U)`` with the
soft-Dice statistics taken over the WHOLE batch.  `CEDiceLoss`, `dice_coeff` and `multiclass_dice_coeff` are the
north-star (upstream milesial UNet) recipe that the reference only documents (model/modelsummary.txt:153-247):
cross-entropy + (1 - Dice) on softmax probabilities, and the Dice score used for evaluation.
These are the torch-op versions (CPU plumbing path, oracles); on a B200 the training step uses the fused
head+loss kernels in csrc/kernels/loss_optim.cu, which compute the same numbers from logits in fp32.
"""
from __future__ import annotations

import torch
import torch.nn.functional as F


class Loss:
    """BCE - log(soft Dice), the reference's criterion. Plain callable like the reference (not an nn.Module).

    As in the reference, `dice_weight` is only a switch: any non-zero value subtracts the un-scaled log-Dice
    term (utils/utils.py:16,23)."""

    def __init__(self, dice_weight: float = 1):
        self.dice_weight = dice_weight

    def __call__(self, outputs: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        outputs = outputs.float()
        targets = targets.float()
        loss = F.binary_cross_entropy(outputs, targets)
        if self.dice_weight:
            hit = (targets == 1).to(outputs.dtype)
            inter = (outputs * hit).sum()
            union = outputs.sum() + hit.sum() + 1e-15
            loss = loss - torch.log(2 * inter / union)
        return loss


def dice_coeff(pred: torch.Tensor, target: torch.Tensor, reduce_batch_first: bool = False,
               epsilon: float = 1e-6) -> torch.Tensor:
    """Dice coefficient of two masks / probability maps of identical shape ([N,H,W] or [H,W])."""
    assert pred.shape == target.shape
    dims = (-1, -2) if pred.dim() == 2 or not reduce_batch_first else (-1, -2, -3)
    inter = 2 * (pred * target).sum(dim=dims)
    sets = pred.sum(dim=dims) + target.sum(dim=dims)
    sets = torch.where(sets == 0, inter, sets)
    return ((inter + epsilon) / (sets + epsilon)).mean()


def multiclass_dice_coeff(pred: torch.Tensor, target: torch.Tensor, reduce_batch_first: bool = False,
                          epsilon: float = 1e-6) -> torch.Tensor:
    """Average Dice over classes; pred/target are [N,C,H,W]."""
    return dice_coeff(pred.flatten(0, 1), target.flatten(0, 1), reduce_batch_first, epsilon)


def dice_loss(pred: torch.Tensor, target: torch.Tensor, multiclass: bool = False) -> torch.Tensor:
    fn = multiclass_dice_coeff if multiclass else dice_coeff
    return 1 - fn(pred, target, reduce_batch_first=True)


class CEDiceLoss:
    """Cross-entropy (or BCE-with-logits for one class) + Dice loss on logits [N,C,H,W]; targets are class ids."""

    def __init__(self, n_classes: int = 2):
        self.n_classes = n_classes

    def __call__(self, logits: torch.Tensor, targets: torch.Tensor) -> torch.Tensor:
        logits = logits.float()
        if self.n_classes == 1:
            t = targets.float().reshape(logits.shape[0], *logits.shape[2:])
            z = logits.squeeze(1)
            return F.binary_cross_entropy_with_logits(z, t) + dice_loss(torch.sigmoid(z), t, multiclass=False)
        t = targets.long().reshape(logits.shape[0], *logits.shape[2:])
        onehot = F.one_hot(t, self.n_classes).permute(0, 3, 1, 2).float()
        return F.cross_entropy(logits, t) + dice_loss(F.softmax(logits, dim=1), onehot, multiclass=True)


@torch.no_grad()
def dice_score(outputs: torch.Tensor, targets: torch.Tensor, n_classes: int = 1) -> torch.Tensor:
    """Evaluation Dice: hard masks from probabilities (n_classes == 1: threshold 0.5) or logits (argmax)."""
    if n_classes == 1:
        pred = (outputs.float().reshape(targets.shape[0], -1) > 0.5).float()
        t = targets.float().reshape(targets.shape[0], -1)
        return dice_coeff(pred.unsqueeze(1), t.unsqueeze(1), reduce_batch_first=False)
    t = targets.long().reshape(outputs.shape[0], *outputs.shape[2:])
    onehot = F.one_hot(t, n_classes).permute(0, 3, 1, 2).float()
    pred = F.one_hot(outputs.argmax(dim=1), n_classes).permute(0, 3, 1, 2).float()
    return multiclass_dice_coeff(pred[:, 1:], onehot[:, 1:], reduce_batch_first=False)
