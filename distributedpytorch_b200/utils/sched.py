"""ReduceLROnPlateau with the semantics the reference relies on (utils/train_utils.py:46,86: mode 'min',
patience 2, torch defaults factor 0.1, relative threshold 1e-4), written against a plain "set_lr" callable so
it can drive the device-resident learning rate of the fused optimizers.  In DDP every rank steps it with the
same (broadcast) validation loss, fixing the reference's rank-0-only step that lets replicas diverge
(utils/train_utils.py:235-241)."""
from __future__ import annotations

from typing import Callable, Optional


class PlateauScheduler:
    def __init__(self, lr: float, set_lr: Optional[Callable[[float], None]] = None, mode: str = "min",
                 factor: float = 0.1, patience: int = 2, threshold: float = 1e-4, min_lr: float = 0.0,
                 eps: float = 1e-8, verbose: bool = False):
        if mode not in ("min", "max"):
            raise ValueError("mode must be 'min' or 'max'")
        self.lr = float(lr)
        self._set_lr = set_lr
        self.mode, self.factor, self.patience = mode, factor, patience
        self.threshold, self.min_lr, self.eps, self.verbose = threshold, min_lr, eps, verbose
        self.best = float("inf") if mode == "min" else -float("inf")
        self.num_bad = 0
        self.last_epoch = 0

    def _better(self, value: float) -> bool:
        if self.mode == "min":
            return value < self.best * (1.0 - self.threshold)
        return value > self.best * (1.0 + self.threshold)

    def step(self, metric: float) -> float:
        value = float(metric)
        self.last_epoch += 1
        if self._better(value):
            self.best = value
            self.num_bad = 0
        else:
            self.num_bad += 1
        if self.num_bad > self.patience:
            new_lr = max(self.lr * self.factor, self.min_lr)
            if self.lr - new_lr > self.eps:
                if self.verbose:
                    print(f"PlateauScheduler: reducing learning rate from {self.lr:.3e} to {new_lr:.3e}")
                self.lr = new_lr
                if self._set_lr is not None:
                    self._set_lr(new_lr)
            self.num_bad = 0
        return self.lr

    def state_dict(self) -> dict:
        return {"lr": self.lr, "best": self.best, "num_bad": self.num_bad, "last_epoch": self.last_epoch}

    def load_state_dict(self, sd: dict) -> None:
        self.lr, self.best = sd["lr"], sd["best"]
        self.num_bad, self.last_epoch = sd["num_bad"], sd["last_epoch"]
        if self._set_lr is not None:
            self._set_lr(self.lr)
