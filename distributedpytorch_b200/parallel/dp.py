"""Single-process multi-GPU data parallelism (`-t DP`) without torch.nn.DataParallel.

Reference: `DP(model, device_ids=range(WORLD_SIZE))` (utils/train_utils.py:98): every forward scatters the batch
from GPU0, re-broadcasts all parameters from GPU0, runs the replicas in Python threads, gathers the outputs on
GPU0 where the loss is computed (whole-batch Dice), and the backward reduce-adds every replica's gradients onto
GPU0, where a root-only optimizer steps.

`NativeDP` keeps those semantics (global batch B, loss statistics over the WHOLE batch, one optimizer update,
identical replicas) and removes the four per-step collectives as separate operations:
  scatter    - each replica's input-packing kernel reads its chunk of the batch straight out of the root GPU's
               memory over NVLink (or its own slice of the pinned host batch); no scatter kernel, no staging copy
  broadcast  - the fused update kernel writes the new fp32 weights into every replica's arena (all-gather half)
  gather     - each replica's head kernel stores its probabilities directly into the root's output tensor; the
               loss partial sums are all-reduced (tiny) so every replica back-propagates the whole-batch loss
  reduce     - the fused update kernel pulls each rank's slice of the gradients from every replica (reduce-scatter
               half) and applies Adam in the same pass (root-only optimizer semantics, sharded execution)
Replica programs are launched concurrently by native worker threads (csrc/runtime/program.cu), not by Python
threads holding the GIL (torch/nn/parallel/parallel_apply.py).

Both UNet variants run here.  For the north-star (BatchNorm) variant every replica normalises with the statistics of
its own chunk and only the root's running statistics survive a step, exactly what torch's DataParallel does (replicas
are re-created from the root module every forward); they are copied to the replicas before an evaluation forward.

`TorchDP` is the same math on plain torch ops on one device (CPU plumbing path / oracle): the replicas share
the root's parameters, each runs its chunk, outputs are concatenated and the loss is taken on the whole batch.
"""
from __future__ import annotations

import copy
import ctypes as C
from typing import List, Optional, Sequence

import torch

from ..models.unet import UNet


def chunk_sizes(batch: int, n: int) -> List[int]:
    """torch.nn.parallel.scatter semantics: chunks of ceil(B/n); with B < n only B devices are used."""
    per = -(-batch // n)
    out, left = [], batch
    while left > 0:
        out.append(min(per, left))
        left -= out[-1]
    return out


class NativeDP:
    def __init__(self, model: UNet, batch: int, height: int, width: int, device_ids: Sequence[int], *, lr: float = 1e-4,
                 weight_decay: float = 1e-8, optimizer: str = "adam", eps: float = 1e-8, betas=(0.9, 0.999),
                 comm_blocks: int = 64, algo: str = "auto", momentum: float = 0.999, alpha: float = 0.99):
        """`algo`: "two_shot" / "one_shot" (peer loads and stores) or "multimem" (the gradient and parameter arenas of all
        replicas live on VMM allocations bound to one NVLS multicast object: the reduce half is `multimem.ld_reduce`, summed
        inside the NVSwitch, and the broadcast half one `multimem.st` per value instead of world-1 peer stores);
        "auto" = multimem from 4 devices up when the switch supports it, like FusedDDP."""
        from ..engine.milesial import NativeMilesialEngine
        from ..engine.native_unet import NativeUNetEngine
        from ..engine.program import run_parallel
        from ..ops import native as N
        from .comm import DeviceTeam

        self._run_parallel = run_parallel
        self.N = N
        sizes = chunk_sizes(batch, len(device_ids))
        self.devices = [int(d) for d in device_ids][:len(sizes)]
        self.sizes = sizes
        self.world = len(self.devices)
        self.B, self.H, self.W = batch, height, width
        self.model = model
        self.team = DeviceTeam(self.devices)
        if algo == "auto":
            algo = "multimem" if self.world >= 4 and self.team.multicast_supported() else "two_shot"
        if algo == "multimem" and not self.team.multicast_supported():
            raise RuntimeError("algo='multimem' needs >= 2 devices on an NVSwitch domain with multicast support")
        n_arena = sum((p.numel() + 3) // 4 * 4 for p in model.parameters())
        self.mc_params = self.team.mc_buffer(n_arena * 4) if algo == "multimem" else None
        self.mc_grads = self.team.mc_buffer(n_arena * 4) if algo == "multimem" else None
        total_pixels = batch * height * width
        root_dev = torch.device("cuda", self.devices[0])
        # root-resident batch and gathered output (what the reference's loop sees on cuda:0)
        self.milesial = model.variant == "milesial"
        self.images = torch.zeros(batch, 3, height, width, device=root_dev)
        self.target = torch.zeros(batch, height, width, device=root_dev, dtype=torch.uint8 if self.milesial else torch.float32)
        # gathered output: probabilities [B,1,H,W] (reference variant) or logits [B,K,H,W] (north-star variant)
        self.prob = torch.zeros(batch, model.n_classes if self.milesial else 1, height, width, device=root_dev)
        self.wants_float_masks = not self.milesial
        offs = [0]
        for b in sizes:
            offs.append(offs[-1] + b)
        hb = N.lib().dpt_head_blocks()
        # loss partial sums of every replica (peer-visible), all-reduced before each replica finalizes the loss
        self.partials = [torch.zeros(hb, 4, device=f"cuda:{d}") for d in self.devices]
        partial_ptrs = (C.c_void_p * self.world)(*[t.data_ptr() for t in self.partials])
        self._keep = [partial_ptrs]
        self.engines: List[NativeUNetEngine] = []
        self.streams: List[torch.cuda.Stream] = []
        for r, (d, b) in enumerate(zip(self.devices, sizes)):
            dev = torch.device("cuda", d)
            replica = model if r == 0 else copy.deepcopy(model)
            io = {"images": self.images[offs[r]:offs[r + 1]], "target": self.target[offs[r]:offs[r + 1]],
                  # scatter / gather become peer loads / stores
                  ("logits" if self.milesial else "prob"): self.prob[offs[r]:offs[r + 1]],
                  "head_partial": self.partials[r]}

            def pre_finalize(prog, r=r):
                prog.allreduce_sum(self.team.structs[r], partial_ptrs, hb * 4, 1.0, 4,
                                   "whole-batch loss statistics all-reduce")

            arenas = {}
            if algo == "multimem":
                arenas = {"params": self.mc_params.tensor(r, torch.float32, n_arena),
                          "grads": self.mc_grads.tensor(r, torch.float32, n_arena)}
            if self.milesial:
                eng = NativeMilesialEngine(replica, b, height, width, dev, lr=lr, weight_decay=weight_decay, optimizer=optimizer,
                                           eps=eps, betas=betas, momentum=momentum, alpha=alpha, loss_scale=float(batch),
                                           total_pixels=total_pixels, build_update=False, io_tensors=io,
                                           pre_finalize=pre_finalize, **arenas)
            else:
                eng = NativeUNetEngine(replica, b, height, width, dev, lr=lr, weight_decay=weight_decay, optimizer=optimizer,
                                       eps=eps, betas=betas, loss_scale=float(batch), total_pixels=total_pixels,
                                       build_update=False, io_tensors=io, pre_finalize=pre_finalize, **arenas)
            self.engines.append(eng)
            with torch.cuda.device(dev):
                self.streams.append(torch.cuda.Stream(device=dev))
        root = self.engines[0]
        if self.milesial:
            n = root.numel
            grads = [e.grads.data_ptr() for e in self.engines]
            params = [e.params.data_ptr() for e in self.engines]
        else:
            n = root.arena.numel
            grads = [e.arena.grads.data_ptr() for e in self.engines]
            params = [e.arena.params.data_ptr() for e in self.engines]
        for r, eng in enumerate(self.engines):
            with torch.cuda.device(eng.device):
                d = N.FusedArDesc()
                for k in range(self.world):
                    d.grads[k], d.params[k] = grads[k], params[k]
                d.state0, d.opt_state = eng.m.data_ptr(), eng.opt_state.data_ptr()
                d.state1 = eng.v.data_ptr() if (optimizer == "adam" or eng.rms_momentum > 0) else None
                d.n = n
                d.algo = {"two_shot": N.AR_TWO_SHOT, "one_shot": N.AR_ONE_SHOT, "multimem": N.AR_MULTIMEM}[algo]
                if algo == "multimem":
                    d.mc_grads, d.mc_params = self.mc_grads.mc_ptrs[r], self.mc_params.mc_ptrs[r]
                d.optimizer = N.OPT_ADAM if optimizer == "adam" else N.OPT_RMSPROP
                d.blocks = comm_blocks
                d.eps, d.weight_decay, d.grad_scale = eps, weight_decay, 1.0  # DataParallel SUMS replica gradients
                d.alpha, d.momentum = eng.rms_alpha, eng.rms_momentum
                prog = eng.fwd
                prog.opt_tick(eng.opt_state)
                prog.fused_allreduce(self.team.structs[r], d, f"fused reduce+{optimizer}+broadcast (device {r})")
                prog.prep_weights(eng.prep_table, eng.prep_entries)
                eng.step_end = len(prog)
        self.programs = [e.fwd for e in self.engines]
        self.loss_state = root.loss_state
        self.algo, self.comm_blocks, self.arena_numel = algo, comm_blocks, n
        # running statistics of the replicas' BatchNorm layers, refreshed from the root before evaluation forwards
        self._bn_buffers = [[b for b in e.model.buffers() if b.dtype == torch.float32] for e in self.engines] if self.milesial else None
        self._bn_stale = False
        self.ready = torch.cuda.Event()
        self.done = [torch.cuda.Event() for _ in self.devices]

    # ------------------------------------------------------------------------------------------------
    def load_batch(self, images: torch.Tensor, masks: torch.Tensor) -> None:
        root = torch.device("cuda", self.devices[0])
        with torch.cuda.device(root):
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
            masks = masks.reshape(self.target.shape)
            self.target.copy_(masks if masks.dtype == self.target.dtype else masks.to(self.target.dtype), non_blocking=True)
            self.ready.record(torch.cuda.current_stream(root))

    def _run(self, first_attr: Optional[str], last_attr: str, programs=None) -> None:
        for s in self.streams:
            s.wait_event(self.ready)
        first = [0] * self.world
        last = [getattr(e, last_attr) if last_attr else -1 for e in self.engines]
        aux = [getattr(e, "aux_stream", None) or s for e, s in zip(self.engines, self.streams)]  # two-stream backward
        self._run_parallel(programs or self.programs, self.streams, aux, first, last)
        root = torch.device("cuda", self.devices[0])
        for ev, s in zip(self.done, self.streams):
            ev.record(s)
        cur = torch.cuda.current_stream(root)
        for ev in self.done:
            cur.wait_event(ev)

    def train_step(self, images=None, masks=None, stream=None) -> torch.Tensor:
        if images is not None:
            self.load_batch(images, masks)
        if self.milesial:
            for e, s in zip(self.engines, self.streams):
                with torch.cuda.device(e.device):
                    e._enter_train(s)
            self._bn_stale = True
        self._run(None, "step_end")
        return self.loss_state[4]

    def _eval_forward(self) -> None:
        """North-star variant: evaluation-mode forward of every replica (root's running statistics, BatchNorm folded)."""
        if self._bn_stale:
            for bufs, e, s in zip(self._bn_buffers[1:], self.engines[1:], self.streams[1:]):
                with torch.cuda.device(e.device), torch.cuda.stream(s):
                    s.wait_event(self.ready)
                    for dst, src in zip(bufs, self._bn_buffers[0]):
                        dst.copy_(src, non_blocking=True)
            self._bn_stale = False
        for e, s in zip(self.engines, self.streams):
            with torch.cuda.device(e.device):
                e._enter_eval(s)
        self._run(None, "", programs=[e.evalp for e in self.engines])

    def evaluate_batch(self, images, masks, stream=None) -> torch.Tensor:
        self.load_batch(images, masks)
        if self.milesial:
            from ..utils.loss import CEDiceLoss

            self._eval_forward()
            with torch.cuda.device(self.prob.device):
                return CEDiceLoss(self.prob.shape[1])(self.prob, self.target.long())
        self._run(None, "fwd_end")
        return self.loss_state[4]

    def forward(self, images, stream=None) -> torch.Tensor:
        """Gathered probabilities [B,1,H,W] (logits [B,K,H,W] for the north-star variant) on the root device, like
        DataParallel.forward."""
        self.load_batch(images, torch.zeros_like(self.target))
        if self.milesial:
            self._eval_forward()
        else:
            self._run(None, "fwd_end")
        return self.prob

    def set_lr(self, lr: float) -> None:
        for e in self.engines:
            with torch.cuda.device(e.device):
                e.set_lr(lr)

    def optimizer_state(self) -> dict:
        """The update is sharded over the replicas (each owns a slice of the moments): their sum is the full state."""
        self.synchronize()
        m = sum(e.m.detach().cpu() for e in self.engines)
        v = sum(e.v.detach().cpu() for e in self.engines)
        return {"m": m, "v": v, "opt_state": self.engines[0].opt_state.detach().cpu()}

    def load_optimizer_state(self, sd: dict) -> None:
        for e in self.engines:
            with torch.cuda.device(e.device):
                e.load_optimizer_state(sd)

    def synchronize(self) -> None:
        for d in self.devices:
            torch.cuda.synchronize(d)


class TorchDP:
    """DataParallel math on one device with plain torch ops (CPU plumbing / oracle)."""

    def __init__(self, model: torch.nn.Module, criterion, n_replicas: int, *, lr: float = 1e-4, weight_decay: float = 1e-8,
                 optimizer: str = "adam", loss_scale: float = 1.0, device=None):
        from ..trainer.steppers import TorchStepper

        self.n = n_replicas
        self.inner = TorchStepper(model, criterion, lr=lr, weight_decay=weight_decay, optimizer=optimizer,
                                  loss_scale=loss_scale, device=device)
        self.model = model
        self.criterion = criterion
        self.device = self.inner.device

    def _gathered(self, images: torch.Tensor) -> torch.Tensor:
        sizes = chunk_sizes(images.shape[0], self.n)
        outs, off = [], 0
        for b in sizes:  # every replica runs its chunk with the root's parameters
            outs.append(self.model(images[off:off + b]))
            off += b
        return torch.cat(outs, dim=0)  # gather on the root

    def train_step(self, images, masks, stream=None):
        images, masks = self.inner._prep(images, masks)
        loss = self.criterion(self._gathered(images), masks)  # loss on the whole gathered batch
        self.inner.opt.zero_grad(set_to_none=True)
        (self.inner.loss_scale * loss).backward()  # gradients of all replicas accumulate on the root parameters
        self.inner.opt.step()
        return loss.detach()

    @torch.no_grad()
    def evaluate_batch(self, images, masks, stream=None):
        images, masks = self.inner._prep(images, masks)
        return self.criterion(self._gathered(images), masks)

    @torch.no_grad()
    def forward(self, images, stream=None):
        return self._gathered(images.to(self.device).float())

    def set_lr(self, lr):
        self.inner.set_lr(lr)

    supports_partial_batch = True

    def optimizer_state(self):
        return self.inner.optimizer_state()

    def load_optimizer_state(self, sd):
        self.inner.load_optimizer_state(sd)
