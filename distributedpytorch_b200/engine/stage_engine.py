"""Native pipeline stage: blocks [lo, hi) of the UNet on one B200, one NativeUNetEngine per in-flight micro-batch
sharing a single parameter core (arena, kernel-side weights, optimizer state).  Implements the StageBackend
interface of parallel/pipe.py; boundary tensors are the engines' own static NHWC bf16 buffers, so NCCL sends and
receives read and write exactly the memory the kernels use (no staging copies, bf16 on the wire: 72.7 MiB per
image per direction for the reference's 2-stage cut instead of 145.3 MiB in fp32)."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import native as N
from .native_unet import NUM_BLOCKS, NativeUNetEngine
from .program import Program


class NativeStage:
    def __init__(self, model, lo: int, hi: int, mb_sizes: Sequence[int], height: int, width: int, device, *,
                 lr: float = 1e-4, weight_decay: float = 1e-8, optimizer: str = "adam", loss_scale: Optional[float] = None,
                 whole_batch: bool = True):
        if len(set(mb_sizes)) != 1:
            raise ValueError("the native pipeline needs equal micro-batches (batch divisible by their number)")
        self.blocks = (lo, hi)
        self.device = torch.device(device)
        self.mb_sizes = list(mb_sizes)
        self.M = len(mb_sizes)
        b = mb_sizes[0]
        total = sum(mb_sizes)
        self.is_last = hi == NUM_BLOCKS
        self.whole_batch = whole_batch
        self.model = model
        scale = float(total if loss_scale is None else loss_scale)
        with torch.cuda.device(self.device):
            self.loss_state = torch.zeros(8, device=self.device)
            self.engines: List[NativeUNetEngine] = []
            for mb in range(self.M):
                io = {"loss_state": self.loss_state} if whole_batch else {}
                eng = NativeUNetEngine(
                    model, b, height, width, self.device, lr=lr, weight_decay=weight_decay, optimizer=optimizer,
                    loss_scale=scale, total_pixels=total * height * width, blocks=(lo, hi),
                    shared=self.engines[0] if mb else None, accumulate=True, build_update=False,
                    finalize=not whole_batch, dice_scale=1.0 if whole_batch else b / total, io_tensors=io)
                self.engines.append(eng)
            core = self.engines[0]
            self.core = core
            self.aux = Program(self.device)
            self._zero = self.aux.memset(core.arena.grads) if core.arena.numel else None
            self.zero_end = len(self.aux)
            if self.is_last and whole_batch:
                self.aux.head_finalize([e.head_partial for e in self.engines], core.head_blocks, core.total_pixels,
                                       core.dice_weight, self.loss_state, None, 0, "whole-batch loss finalize")
            self.fin_end = len(self.aux)
            core._build_update(self.aux)
            self.upd_end = len(self.aux)
            self._loss_acc = torch.zeros((), device=self.device)

    # ---- StageBackend ------------------------------------------------------------------------------------
    def input_buffer(self, flow: str, mb: int) -> torch.Tensor:
        return self.engines[mb].act[NativeUNetEngine.flow_key(flow)]

    def output_tensor(self, flow: str, mb: int) -> torch.Tensor:
        return self.engines[mb].act[NativeUNetEngine.flow_key(flow)]

    def grad_in_buffer(self, flow: str, mb: int) -> torch.Tensor:
        return self.engines[mb].grad[NativeUNetEngine.flow_grad_key(flow)]

    def grad_out_tensor(self, flow: str, mb: int) -> torch.Tensor:
        return self.engines[mb].grad[NativeUNetEngine.flow_grad_key(flow)]

    def load_micro_batch(self, mb: int, images, masks) -> None:
        with torch.cuda.device(self.device):
            self.engines[mb].load_batch(images if self.blocks[0] == 0 else None, masks if self.is_last else None)

    def begin_step(self) -> None:
        if self.zero_end:
            self.aux.run(0, self.zero_end)
        if self.is_last and not self.whole_batch:
            self._loss_acc.zero_()

    def forward(self, mb: int) -> None:
        self.engines[mb].run_forward()

    def finalize_loss(self, mbs: Sequence[int]) -> None:
        if self.whole_batch:
            self.aux.run(self.zero_end, self.fin_end)
        else:
            for mb in mbs:
                self._loss_acc += self.engines[mb].loss_state[4]

    def backward(self, mb: int) -> None:
        self.engines[mb].run_backward()

    def optimizer_step(self) -> None:
        self.aux.run(self.fin_end, self.upd_end)

    def loss(self) -> torch.Tensor:
        return self.loss_state[4] if self.whole_batch else self._loss_acc

    def set_lr(self, lr: float) -> None:
        with torch.cuda.device(self.device):
            self.core.set_lr(lr)

    def probabilities(self) -> torch.Tensor:
        return torch.cat([e.prob for e in self.engines], dim=0)
