"""Native pipeline stage: blocks [lo, hi) of the UNet on one B200, one engine (NativeUNetEngine for the reference
variant, NativeMilesialEngine for the north-star BatchNorm variant) per in-flight micro-batch
sharing a single parameter core (arena, kernel-side weights, optimizer state).  Implements the StageBackend
interface of parallel/pipe.py; boundary tensors are the engines' own static NHWC bf16 buffers, so NCCL sends and
receives read and write exactly the memory the kernels use (no staging copies, bf16 on the wire: 72.7 MiB per
image per direction for the reference's 2-stage cut instead of 145.3 MiB in fp32)."""
from __future__ import annotations

from typing import List, Optional, Sequence

import torch

from ..ops import native as N
from .milesial import NativeMilesialEngine
from .native_unet import NUM_BLOCKS, NativeUNetEngine
from .program import Program


class NativeStage:
    def __init__(self, model, lo: int, hi: int, mb_sizes: Sequence[int], height: int, width: int, device, *,
                 lr: float = 1e-4, weight_decay: float = 1e-8, optimizer: Optional[str] = None, loss_scale: Optional[float] = None,
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
        self.milesial = model.variant == "milesial"
        if self.milesial:
            # The softmax head's gradient depends only on the pixel count of the whole batch (the Dice denominator is a
            # constant), so every micro-batch back-propagates the exact whole-batch loss under GPipe AND 1F1B; the loss
            # VALUE is summed over the micro-batches' partial sums at the end of the step.  BatchNorm: per-micro-batch
            # statistics, running statistics updated once per micro-batch.
            self.whole_batch = whole_batch = True
            self._deferred_loss = True
        with torch.cuda.device(self.device):
            self.loss_state = torch.zeros(8, device=self.device)
            self.engines: List = []
            for mb in range(self.M):
                io = {"loss_state": self.loss_state} if whole_batch else {}
                if self.milesial:
                    eng = NativeMilesialEngine(
                        model, b, height, width, self.device, lr=lr, weight_decay=weight_decay,
                        optimizer="rmsprop" if optimizer in (None, "rmsprop") else optimizer, loss_scale=scale,
                        total_pixels=total * height * width, blocks=(lo, hi), shared=self.engines[0] if mb else None,
                        accumulate=True, build_update=False, finalize=False, io_tensors=io)
                else:
                    eng = NativeUNetEngine(
                        model, b, height, width, self.device, lr=lr, weight_decay=weight_decay, optimizer=optimizer or "adam",
                        loss_scale=scale, total_pixels=total * height * width, blocks=(lo, hi),
                        shared=self.engines[0] if mb else None, accumulate=True, build_update=False,
                        finalize=not whole_batch, dice_scale=1.0 if whole_batch else b / total, io_tensors=io)
                self.engines.append(eng)
            core = self.engines[0]
            self.core = core
            self.aux = Program(self.device)
            grads, numel = (core.grads, core.numel) if self.milesial else (core.arena.grads, core.arena.numel)
            self.numel = numel
            self._zero = self.aux.memset(grads) if numel else None
            self.zero_end = len(self.aux)
            if self.is_last and whole_batch:
                if self.milesial:
                    self.aux.head_softmax_finalize([e.head_partial for e in self.engines], core.total_pixels, self.loss_state,
                                                   "whole-batch loss finalize")
                else:
                    self.aux.head_finalize([e.head_partial for e in self.engines], core.head_blocks, core.total_pixels,
                                           core.dice_weight, self.loss_state, None, 0, "whole-batch loss finalize")
            self.fin_end = len(self.aux)
            core._build_update(self.aux)
            self.upd_end = len(self.aux)
            self._loss_acc = torch.zeros((), device=self.device)

    # ---- StageBackend ------------------------------------------------------------------------------------
    def input_buffer(self, flow: str, mb: int) -> torch.Tensor:
        e = self.engines[mb]
        return e.act[e.flow_key(flow)]

    def output_tensor(self, flow: str, mb: int) -> torch.Tensor:
        e = self.engines[mb]
        return e.act[e.flow_key(flow)]

    def grad_in_buffer(self, flow: str, mb: int) -> torch.Tensor:
        e = self.engines[mb]
        return e.grad[e.flow_grad_key(flow)]

    def grad_out_tensor(self, flow: str, mb: int) -> torch.Tensor:
        e = self.engines[mb]
        return e.grad[e.flow_grad_key(flow)]

    def load_micro_batch(self, mb: int, images, masks) -> None:
        with torch.cuda.device(self.device):
            self.engines[mb].load_batch(images if self.blocks[0] == 0 else None, masks if self.is_last else None)

    def begin_step(self) -> None:
        if self.zero_end:
            self.aux.run(0, self.zero_end)
        if self.is_last and not self.whole_batch:
            self._loss_acc.zero_()

    block_granular = True  # PipelineStage may run, wait and send block by block

    def forward(self, mb: int) -> None:
        self.engines[mb].run_forward()

    def forward_block(self, mb: int, block: int, evaluate: bool = False) -> None:
        e = self.engines[mb]
        if self.milesial:
            if evaluate:
                e._enter_eval()
                lo, hi = e.infer_block_range[block]
                e.evalp.run(lo, hi)
                return
            e._enter_train()
        lo, hi = e.fwd_block_range[block]
        e.fwd.run(lo, hi)

    def backward_block(self, mb: int, block: int) -> None:
        e = self.engines[mb]
        lo, hi = e.bwd_block_range[block]
        e.fwd.run(lo, hi)

    def eval_forward(self, mb: int) -> None:
        """Forward in evaluation mode (north-star variant: running statistics, BatchNorm folded into the weights)."""
        if self.milesial:
            self.engines[mb].run_eval_forward()
        else:
            self.engines[mb].run_forward()

    def finalize_loss(self, mbs: Sequence[int]) -> None:
        if getattr(self, "_deferred_loss", False):
            if len(mbs) == self.M or mbs[-1] == self.M - 1:  # all forwards done: sum every micro-batch's partials
                self.aux.run(self.zero_end, self.fin_end)
        elif self.whole_batch:
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

    def eval_loss(self) -> torch.Tensor:
        """Validation loss after forward_only(evaluate=True) (last stage)."""
        if not self.milesial:
            return self.loss()
        from ..utils.loss import CEDiceLoss

        logits = torch.cat([e.logits for e in self.engines], dim=0)
        target = torch.cat([e.target for e in self.engines], dim=0)
        return CEDiceLoss(logits.shape[1])(logits, target.long())

    def set_lr(self, lr: float) -> None:
        with torch.cuda.device(self.device):
            self.core.set_lr(lr)

    def probabilities(self) -> torch.Tensor:
        return torch.cat([e.logits if self.milesial else e.prob for e in self.engines], dim=0)

    def owned_parameter_names(self):
        return set(self.core.slots) if self.milesial else set(self.core.arena.slots)
