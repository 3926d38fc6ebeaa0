"""Loss curves in the reference's format (utils/train_utils.py:50-52,75-84,89-92): rows
[global_step, seconds_since_start, loss] pickled as pandas DataFrames with columns Step, Time, Loss under
./loss/<mode>/.  The directories are created here (the reference never creates them)."""
from __future__ import annotations

import os
import time
from typing import List, Optional

import numpy as np


class LossRecorder:
    def __init__(self, mode: str, directory: str = "./loss", every: int = 10):
        self.mode, self.directory, self.every = mode, directory, every
        self.train_rows: List[list] = []
        self.val_rows: List[list] = []
        self._recent: List[float] = []
        self.start_time = time.time()

    def elapsed(self) -> float:
        return time.time() - self.start_time

    def train_step(self, global_step: int, loss: float) -> None:
        """Call once per step with the 1-based global step; records the mean of the last `every` losses."""
        self._recent.append(float(loss))
        if global_step % self.every == 0:
            self.train_rows.append([global_step, self.elapsed(), float(np.mean(self._recent[-self.every:]))])
            self._recent = self._recent[-self.every:]

    def validation(self, global_step: int, loss: float) -> None:
        self.val_rows.append([global_step, self.elapsed(), float(loss)])

    def save(self) -> Optional[str]:
        import pandas as pd

        out = os.path.join(self.directory, self.mode)
        os.makedirs(out, exist_ok=True)
        pd.DataFrame(self.train_rows, columns=["Step", "Time", "Loss"]).to_pickle(os.path.join(out, "train_loss.pkl"))
        pd.DataFrame(self.val_rows, columns=["Step", "Time", "Loss"]).to_pickle(os.path.join(out, "val_loss.pkl"))
        return out
