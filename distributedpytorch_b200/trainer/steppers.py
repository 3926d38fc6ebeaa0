"""Step executors with one interface: train_step / evaluate_batch / forward / set_lr.

`TorchStepper` is the torch-op path (CPU plumbing, non-reference variants without native kernels yet, oracle);
`make_stepper` picks the native engine on a CUDA device for the reference variant."""
from __future__ import annotations

from typing import Optional

import torch


class TorchStepper:
    def __init__(self, model: torch.nn.Module, criterion, *, lr: float = 1e-4, weight_decay: float = 1e-8,
                 optimizer: str = "adam", loss_scale: float = 1.0, device: Optional[torch.device] = None,
                 momentum: float = 0.999):
        self.model = model
        self.criterion = criterion
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.loss_scale = loss_scale
        params = [p for p in model.parameters() if p.requires_grad]
        if optimizer == "adam":
            self.opt = torch.optim.Adam(params, lr=lr, weight_decay=weight_decay)
        elif optimizer == "rmsprop":
            self.opt = torch.optim.RMSprop(params, lr=lr, weight_decay=weight_decay, momentum=momentum)
        else:
            raise ValueError(f"unknown optimizer {optimizer!r}")

    def _prep(self, images, masks):
        images = images.to(self.device, non_blocking=True).float()
        masks = masks.to(self.device, non_blocking=True)
        return images, masks

    def train_step(self, images, masks, stream=None) -> torch.Tensor:
        images, masks = self._prep(images, masks)
        loss = self.criterion(self.model(images), masks)
        self.opt.zero_grad(set_to_none=True)
        (self.loss_scale * loss).backward()
        self.opt.step()
        return loss.detach()

    @torch.no_grad()
    def evaluate_batch(self, images, masks, stream=None) -> torch.Tensor:
        images, masks = self._prep(images, masks)
        return self.criterion(self.model(images), masks)

    @torch.no_grad()
    def forward(self, images, stream=None) -> torch.Tensor:
        return self.model(images.to(self.device).float())

    def set_lr(self, lr: float) -> None:
        for g in self.opt.param_groups:
            g["lr"] = lr

    supports_partial_batch = True  # plain torch ops take any batch size

    def optimizer_state(self) -> dict:
        return self.opt.state_dict()

    def load_optimizer_state(self, sd: dict) -> None:
        lr = [g["lr"] for g in self.opt.param_groups]
        self.opt.load_state_dict(sd)
        for g, v in zip(self.opt.param_groups, lr):  # the scheduler owns the learning rate
            g["lr"] = v


class MaskAdapter:
    """Criterion wrapper: the data pipeline yields int64 masks [B,H,W]; the reference criterion wants float
    [B,1,H,W] (utils/train_utils.py:61 `.to(torch.float32).unsqueeze(1)`), the CE+Dice criterion wants class ids."""

    def __init__(self, criterion, binary_prob: bool):
        self.criterion = criterion
        self.binary_prob = binary_prob

    def __call__(self, outputs, masks):
        if self.binary_prob:
            return self.criterion(outputs, masks.float().reshape(outputs.shape))
        return self.criterion(outputs, masks)


def native_available(model, device) -> bool:
    from ..ops import native

    variant = getattr(model, "variant", "")
    ok = variant == "ref" or (variant == "milesial" and getattr(model, "n_classes", 0) >= 2 and getattr(model, "n_channels", 3) <= 16)
    return torch.device(device).type == "cuda" and torch.cuda.is_available() and ok and native.available()


def make_stepper(model, criterion, batch_size: int, size_wh, device, *, lr: float, weight_decay: float = 1e-8,
                 optimizer: str = "adam", force_torch: bool = False, uint8_input: bool = False):
    """Native engine when possible (B200 + reference variant), torch-op path otherwise."""
    device = torch.device(device)
    if not force_torch and native_available(model, device):
        from ..engine import NativeUNetEngine

        w, h = int(size_wh[0]), int(size_wh[1])
        if model.variant == "milesial":
            from ..engine.milesial import NativeMilesialEngine

            return NativeMilesialEngine(model, batch_size, h, w, device, lr=lr, weight_decay=weight_decay,
                                        optimizer=optimizer, loss_scale=float(batch_size)), True
        dt = "uint8" if uint8_input else "float32"  # raw pixels + uint8 masks: normalised on the device
        return NativeUNetEngine(model, batch_size, h, w, device, lr=lr, weight_decay=weight_decay, optimizer=optimizer,
                                dice_weight=getattr(criterion, "dice_weight", 1), input_dtype=dt, target_dtype=dt), True
    on_gpu = device.type == "cuda" and torch.cuda.is_available() and not force_torch
    native_family = getattr(model, "variant", "") == "ref" or (getattr(model, "variant", "") == "milesial"
                                                               and getattr(model, "n_classes", 1) >= 2)
    if on_gpu and native_family:
        from ..ops import native

        native.lib()  # raises: the native library is mandatory on a GPU box, no silent fallback to torch kernels
    model.to(device)
    crit = MaskAdapter(criterion, binary_prob=not getattr(model, "returns_logits", False))
    return TorchStepper(model, crit, lr=lr, weight_decay=weight_decay, optimizer=optimizer, loss_scale=batch_size,
                        device=device), False
