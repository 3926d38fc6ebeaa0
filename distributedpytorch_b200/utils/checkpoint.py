"""Checkpoints in the reference's format: a torch.save'd state_dict of fp32 tensors under ./checkpoints/
(utils/train_utils.py:88,163,244), loadable by the reference's UNet and vice versa.

Differences from the reference, all deliberate: paths are joined portably (train.py:43 uses a
backslash), `module.`-prefixed DP/DDP checkpoints load into a bare model and the other way round, MP-mode
tensors are moved to CPU before saving, and an optional side file `<name>.resume.pth` carries optimizer /
scheduler / step state so training can actually resume (the reference saves weights only)."""
from __future__ import annotations

import os
from collections import OrderedDict
from typing import Any, Dict, Optional

import torch

CHECKPOINT_DIR = "./checkpoints"


def checkpoint_path(name: str, directory: str = CHECKPOINT_DIR) -> str:
    base = name if name.endswith(".pth") else f"{name}.pth"
    return os.path.join(directory, base)


def strip_module_prefix(sd: Dict[str, torch.Tensor]) -> "OrderedDict[str, torch.Tensor]":
    return OrderedDict((k[len("module."):] if k.startswith("module.") else k, v) for k, v in sd.items())


def add_module_prefix(sd: Dict[str, torch.Tensor]) -> "OrderedDict[str, torch.Tensor]":
    return OrderedDict((k if k.startswith("module.") else "module." + k, v) for k, v in sd.items())


def save_checkpoint(model_or_state, name: str, directory: str = CHECKPOINT_DIR, module_prefix: bool = False,
                    resume_state: Optional[Dict[str, Any]] = None) -> str:
    """Save weights as `<directory>/<name>.pth`. `module_prefix=True` reproduces the key names of the
    reference's DP/DDP checkpoints (it saves the wrapper's state_dict)."""
    sd = model_or_state.state_dict() if hasattr(model_or_state, "state_dict") else model_or_state
    sd = OrderedDict((k, v.detach().to("cpu", torch.float32) if v.is_floating_point() else v.detach().cpu())
                     for k, v in sd.items())
    sd = add_module_prefix(sd) if module_prefix else strip_module_prefix(sd)
    os.makedirs(directory, exist_ok=True)
    path = checkpoint_path(name, directory)
    torch.save(sd, path)
    if resume_state is not None:
        torch.save(resume_state, path[:-4] + ".resume.pth")
    return path


def load_checkpoint(model: torch.nn.Module, name: str, directory: str = CHECKPOINT_DIR, strict: bool = True) -> None:
    """Load `<directory>/<name>.pth` (or an explicit file path) into `model`, with or without `module.` keys."""
    path = name if os.path.isfile(name) else checkpoint_path(name, directory)
    sd = torch.load(path, map_location="cpu", weights_only=True)
    model.load_state_dict(strip_module_prefix(sd), strict=strict)


def load_resume_state(name: str, directory: str = CHECKPOINT_DIR) -> Optional[Dict[str, Any]]:
    path = checkpoint_path(name, directory)[:-4] + ".resume.pth"
    return torch.load(path, map_location="cpu", weights_only=False) if os.path.isfile(path) else None
