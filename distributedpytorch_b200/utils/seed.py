"""Seeding / determinism (reference: utils/utils.py:28-35 `set_seed`, called at train.py:35)."""
from __future__ import annotations

import os
import random

import numpy as np
import torch


def set_seed(seed: int, deterministic: bool = True) -> None:
    """Seed python, numpy and torch (CPU + every CUDA device) and export PYTHONHASHSEED.

    The reference sets ``cudnn.deterministic=True`` together with ``cudnn.benchmark=True`` (utils/utils.py:34-35),
    which contradict each other; here `deterministic` selects one behaviour.  The native kernels are
    bit-reproducible by construction (fixed-order split-K and collective reductions), so this only matters for
    the torch-op fallback path.
    """
    random.seed(seed)
    os.environ["PYTHONHASHSEED"] = str(seed)
    np.random.seed(seed % (2**32))
    torch.manual_seed(seed)
    if torch.cuda.is_available():
        torch.cuda.manual_seed_all(seed)
    torch.backends.cudnn.deterministic = bool(deterministic)
    torch.backends.cudnn.benchmark = not deterministic
