"""DistributedDataParallel without torch's DistributedDataParallel.

Reference: `DDP(model, device_ids=[rank])` + `optim.Adam(lr * world_size)` (utils/train_utils.py:195-199): torch's
C++ Reducer buckets gradients, all-reduces them with NCCL during backward, and a separate multi-tensor Adam
then updates every replica.

`FusedDDP` (GPU): the parameter and gradient arenas of every rank live in symmetric memory; after backward ONE
kernel per bucket does reduce-scatter (P2P loads from every peer) + 1/world scaling + Adam on the rank's slice of
the fp32 master weights + all-gather of the updated weights (P2P stores into every peer's arena).  The rank-0
parameter broadcast of DDP's constructor is the same machinery (`dpt_team_bcast`).  Optimizer state is sharded
across ranks as a by-product (each rank only ever touches its slice of exp_avg / exp_avg_sq).

`TorchDDP` (CPU / gloo plumbing path and oracle): flat-gradient all_reduce through torch.distributed + a torch
optimizer, same interface.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional

import torch
import torch.distributed as dist

from ..models.unet import UNet


def arena_numel(model: UNet) -> int:
    return sum((p.numel() + 3) // 4 * 4 for p in model.parameters())


class FusedDDP:
    """`model` may be either UNet family member: the reference variant runs on NativeUNetEngine, the north-star
    (BatchNorm) variant on NativeMilesialEngine, whose running statistics are additionally re-synchronised from rank 0
    after every step (torch DDP's broadcast_buffers=True semantics; batch statistics stay per rank, as in torch)."""

    def __init__(self, model: UNet, batch: int, height: int, width: int, device, *, lr: float = 1e-4,
                 weight_decay: float = 1e-8, optimizer: Optional[str] = None, eps: float = 1e-8, betas=(0.9, 0.999),
                 group: Optional[dist.ProcessGroup] = None, algo: str = "auto", comm_blocks: int = 128,
                 target_dtype: str = "float32", loss_scale: Optional[float] = None, overlap: bool = False,
                 input_dtype: str = "float32"):
        """`overlap=True` (reference variant): the arena is cut into three buckets in backward-completion order
        (decoder tail | bottleneck | encoder) and each bucket's fused collective is launched on the communication
        stream as soon as its last weight gradient has been reduced, so it runs under the rest of backward."""
        from ..ops import native as N
        from .comm import ProcessTeam

        self.device = torch.device(device)
        self.team = ProcessTeam(self.device, group)
        self.rank, self.world = self.team.rank, self.team.world
        n = arena_numel(model)
        if algo == "auto":
            # measured on 8 B200 (profiles/collective_8gpu_r1.jsonl, 31 MB arena): NVLS multimem 0.101 ms, two-shot P2P
            # 0.125 ms, NCCL all-reduce + fused Adam 0.245 ms; on 2 GPUs two-shot wins (0.080 vs 0.108 ms)
            from .comm import multicast_supported

            if n * 4 <= 512 * 1024:
                algo = "one_shot"
            elif self.world >= 4 and not overlap and multicast_supported(self.device):
                algo = "multimem"
            else:
                algo = "two_shot"
        if algo == "multimem":  # NVLS: arenas on VMM allocations bound to multicast objects
            from .comm import multicast_supported

            if self.world < 2 or not multicast_supported(self.device):
                raise RuntimeError("algo='multimem' needs >= 2 ranks on an NVSwitch domain with multicast support")
            self.sym_params = self.team.mc_buffer(n * 4)
            self.sym_grads = self.team.mc_buffer(n * 4)
        else:
            self.sym_params = self.team.buffer(n * 4)
            self.sym_grads = self.team.buffer(n * 4)
        params, grads = self.sym_params.tensor(torch.float32, n), self.sym_grads.tensor(torch.float32, n)
        self.bn_sym = None
        if model.variant == "ref":
            from ..engine.native_unet import NativeUNetEngine

            optimizer = optimizer or "adam"
            self._bucket_ops = []
            hook = self._make_bucket_hook(model, optimizer, eps, weight_decay, comm_blocks) if overlap and self.world > 1 else None
            self.engine = NativeUNetEngine(model, batch, height, width, self.device, lr=lr, weight_decay=weight_decay,
                                           optimizer=optimizer, eps=eps, betas=betas, params=params, grads=grads,
                                           build_update=False, target_dtype=target_dtype, loss_scale=loss_scale,
                                           grads_ready_hook=hook, input_dtype=input_dtype)
            e = self.engine
            state0, state1 = e.m, e.v
            alpha, momentum = e.rms_alpha, e.rms_momentum
        else:
            from ..engine.milesial import NativeMilesialEngine

            optimizer = optimizer or "rmsprop"
            # running statistics in one symmetric buffer so rank 0's copy can be pulled by every rank after each step
            bufs = [(nm, b) for nm, b in model.named_buffers() if b.dtype == torch.float32]
            nb = sum((b.numel() + 3) // 4 * 4 for _, b in bufs)
            # (on a multicast object when the arenas are: rank 0 then streams them to everybody with multimem.st)
            self.bn_sym = self.team.mc_buffer(max(nb, 4) * 4) if algo == "multimem" else self.team.buffer(max(nb, 4) * 4)
            flat = self.bn_sym.tensor(torch.float32, max(nb, 4))
            off = 0
            for _, b in bufs:
                view = flat[off:off + b.numel()].view(b.shape)
                view.copy_(b.detach().to(self.device))
                b.data = view
                off += (b.numel() + 3) // 4 * 4
            self.bn_numel = max(nb, 4)
            self.engine = NativeMilesialEngine(model, batch, height, width, self.device, lr=lr, weight_decay=weight_decay,
                                               optimizer=optimizer, eps=eps, betas=betas, params=params, grads=grads,
                                               build_update=False, loss_scale=1.0 if loss_scale is None else loss_scale)
            e = self.engine
            state0, state1 = e.s0, e.s1
            alpha, momentum = e.alpha, e.momentum
        L = N.lib()
        # DDP constructor semantics: every replica starts from rank 0's parameters (utils/train_utils.py:196)
        self._bufs = (C.c_void_p * self.world)(*self.sym_params.ptrs)
        with torch.cuda.device(self.device):
            if algo == "multimem":  # one store stream from rank 0 through the NVSwitch instead of world-1 pulls
                N.check(L.dpt_team_bcast_multimem(C.byref(self.team.struct), self.sym_params.ptrs[0], self.sym_params.mc_ptr, n, 0,
                                                  comm_blocks, N.stream_ptr()))
            else:
                N.check(L.dpt_team_bcast(C.byref(self.team.struct), self._bufs, n, 0, comm_blocks, N.stream_ptr()))
            torch.cuda.synchronize()
            e.refresh_weights()
        d = N.FusedArDesc()
        for r in range(self.world):
            d.grads[r] = self.sym_grads.ptrs[r]
            d.params[r] = self.sym_params.ptrs[r]
        d.state0, d.state1, d.opt_state = state0.data_ptr(), state1.data_ptr(), e.opt_state.data_ptr()
        d.n = n
        d.algo = {"two_shot": N.AR_TWO_SHOT, "one_shot": N.AR_ONE_SHOT, "multimem": N.AR_MULTIMEM}[algo]
        if algo == "multimem":
            d.mc_grads, d.mc_params = self.sym_grads.mc_ptr, self.sym_params.mc_ptr
        d.optimizer = N.OPT_ADAM if optimizer == "adam" else N.OPT_RMSPROP
        d.blocks = comm_blocks
        d.eps, d.weight_decay = eps, weight_decay
        d.grad_scale = 1.0 / self.world  # DDP averages gradients
        d.alpha, d.momentum = alpha, momentum
        self.desc = d
        self.algo = algo
        prog = e.fwd
        if getattr(self, "_bucket_ops", None):
            # buckets were launched on the aux stream during backward; join them before the weights are re-laid out
            for bd in self._bucket_descs:
                bd.state0, bd.state1, bd.opt_state = (state0.data_ptr() + 4 * bd._off, state1.data_ptr() + 4 * bd._off,
                                                      e.opt_state.data_ptr())
            prog.record(7, stream=1, label="buckets done")
            prog.wait(7, stream=0, label="join communication stream")
            with torch.cuda.device(self.device):
                e.aux_stream = torch.cuda.Stream(device=self.device)
            algo = "two_shot x3 buckets, overlapped with backward"
        else:
            prog.opt_tick(e.opt_state)
            prog.fused_allreduce(self.team.struct, d, f"fused allreduce+{optimizer} ({algo}, world {self.world})")
        self.algo = algo
        prog.prep_weights(e.prep_table, e.prep_entries)
        if self.bn_sym is not None and self.world > 1:
            self._bn_bufs = (C.c_void_p * self.world)(*self.bn_sym.ptrs)
            if getattr(self.bn_sym, "mc_ptr", None):
                prog.bcast_multimem(self.team.struct, self.bn_sym.ptrs[0], self.bn_sym.mc_ptr, self.bn_numel, 0, 8,
                                    "BatchNorm running statistics <- rank 0 (multimem.st)")
            else:
                prog.bcast(self.team.struct, self._bn_bufs, self.bn_numel, 0, 8, "BatchNorm running statistics <- rank 0")
        e.step_end = len(prog)
        # expose the engine's step interface
        self.loss_state = e.loss_state
        self.model = model
        self.B = batch
        self.input_uint8 = getattr(e, "input_uint8", False)
        self.wants_float_masks = getattr(e, "wants_float_masks", True)

    # ---- bucketed overlap --------------------------------------------------------------------------------
    BUCKET_TRIGGERS = ("decoder.deconv2", "mid.conv_block.0", "encoder.conv1.conv_block.0")

    def _make_bucket_hook(self, model, optimizer, eps, weight_decay, comm_blocks):
        from ..engine.native_unet import _forward_order_names
        from ..ops import native as N

        have = dict(model.named_parameters())
        offs, off = {}, 0
        for nm in _forward_order_names(model):
            offs[nm] = off
            off += (have[nm].numel() + 3) // 4 * 4
        total = off
        starts = sorted(offs[t + ".weight"] for t in self.BUCKET_TRIGGERS)
        bounds = starts + [total]
        self._bucket_descs = []
        by_trigger = {}
        for k, trig in enumerate(sorted(self.BUCKET_TRIGGERS, key=lambda t: offs[t + ".weight"])):
            lo, hi = bounds[k], bounds[k + 1]
            d = N.FusedArDesc()
            for r in range(self.world):
                d.grads[r] = self.sym_grads.ptrs[r] + 4 * lo
                d.params[r] = self.sym_params.ptrs[r] + 4 * lo
            d.n, d.algo = hi - lo, N.AR_TWO_SHOT
            d.optimizer = N.OPT_ADAM if optimizer == "adam" else N.OPT_RMSPROP
            d.blocks = min(comm_blocks, 32)  # few blocks: the collective shares the GPU with backward kernels
            d.eps, d.weight_decay, d.grad_scale = eps, weight_decay, 1.0 / self.world
            d.alpha, d.momentum = 0.99, 0.0
            d._off = lo
            self._bucket_descs.append(d)
            by_trigger[trig] = (k, d)

        def hook(engine, prog, name):
            if name == "segmap":
                prog.opt_tick(engine.opt_state)
            if name in by_trigger:
                k, d = by_trigger[name]
                prog.record(k, stream=0, label=f"bucket {k} gradients ready")
                prog.wait(k, stream=1, label=f"bucket {k}: communication stream waits")
                prog.fused_allreduce(self.team.struct, d, f"fused allreduce+{optimizer} bucket {k} [{d._off}:{d._off + d.n}]", stream=1)
                self._bucket_ops.append(k)

        return hook

    def load_batch(self, images, masks):
        self.engine.load_batch(images, masks)

    def prefetch(self, images, masks):
        self.engine.prefetch(images, masks)

    def train_step(self, images=None, masks=None, stream=None):
        return self.engine.train_step(images, masks, stream)

    def evaluate_batch(self, images, masks, stream=None):
        return self.engine.evaluate_batch(images, masks, stream)

    def forward(self, images, stream=None):
        return self.engine.forward(images, stream)

    def set_lr(self, lr: float) -> None:
        self.engine.set_lr(lr)

    @property
    def supports_partial_batch(self) -> bool:
        return bool(getattr(self.engine, "supports_partial_batch", False))

    def optimizer_state(self) -> dict:
        """Full optimizer state (COLLECTIVE: every rank calls it).  The fused kernel shards the update, so each rank's
        moment arrays are non-zero only on the slice it owns; their sum over ranks is the complete state."""
        e = self.engine
        out = {}
        for k, t in (("m", e.m), ("v", e.v)):
            full = t.detach().clone()
            if self.world > 1:
                dist.all_reduce(full, group=self.team.group)
            out[k] = full.cpu()
        out["opt_state"] = e.opt_state.detach().cpu()
        return out

    def load_optimizer_state(self, sd: dict) -> None:
        self.engine.load_optimizer_state(sd)  # entries outside the rank's slice are never read

    @property
    def launches_per_step(self) -> int:
        return self.engine.launches_per_step


class TorchDDP:
    """Same semantics on plain torch ops + torch.distributed collectives (works on CPU with gloo)."""

    def __init__(self, model: torch.nn.Module, criterion, *, lr: float = 1e-4, weight_decay: float = 1e-8,
                 optimizer: str = "adam", loss_scale: float = 1.0, group: Optional[dist.ProcessGroup] = None,
                 device: Optional[torch.device] = None):
        self.model = model
        self.criterion = criterion
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.device = torch.device(device) if device is not None else next(model.parameters()).device
        self.loss_scale = loss_scale
        if self.world > 1:  # rank-0 broadcast of module state (parameters AND buffers), like DDP's constructor
            for t in list(model.parameters()) + list(model.buffers()):
                dist.broadcast(t.data, src=dist.get_global_rank(group, 0) if group is not None else 0, group=group)
        params = [p for p in model.parameters() if p.requires_grad]
        self.params = params
        if optimizer == "adam":
            self.opt = torch.optim.Adam(params, lr=lr, weight_decay=weight_decay)
        else:
            self.opt = torch.optim.RMSprop(params, lr=lr, weight_decay=weight_decay, momentum=0.999, foreach=True)
        self.loss_state = torch.zeros(8)

    def _prep(self, images, masks):
        images = images.to(self.device, non_blocking=True).float()
        masks = masks.to(self.device, non_blocking=True)
        return images, masks

    def train_step(self, images, masks, stream=None):
        images, masks = self._prep(images, masks)
        out = self.model(images)
        loss = self.criterion(out, masks)
        self.opt.zero_grad(set_to_none=True)
        (self.loss_scale * loss).backward()
        if self.world > 1:
            flat = torch.cat([p.grad.reshape(-1) for p in self.params])
            dist.all_reduce(flat, group=self.group)
            flat /= self.world
            off = 0
            for p in self.params:
                n = p.numel()
                p.grad.copy_(flat[off:off + n].view_as(p))
                off += n
        self.opt.step()
        return loss.detach()

    @torch.no_grad()
    def evaluate_batch(self, images, masks, stream=None):
        images, masks = self._prep(images, masks)
        return self.criterion(self.model(images), masks)

    @torch.no_grad()
    def forward(self, images, stream=None):
        return self.model(images.to(self.device).float())

    def set_lr(self, lr: float) -> None:
        for g in self.opt.param_groups:
            g["lr"] = lr

    supports_partial_batch = True

    def optimizer_state(self) -> dict:
        return self.opt.state_dict()

    def load_optimizer_state(self, sd: dict) -> None:
        lr = [g["lr"] for g in self.opt.param_groups]
        self.opt.load_state_dict(sd)
        for g, v in zip(self.opt.param_groups, lr):
            g["lr"] = v
