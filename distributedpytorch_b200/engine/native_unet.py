"""Native execution engine for the reference-variant UNet on one B200.

What the reference runs through nn.Module.forward + autograd + torch.optim (utils/train_utils.py:59-70) is
compiled here, once, into a native launch program (engine/program.py -> csrc/runtime/program.cu): forward,
fused head + loss, a hand-written backward pass and the optimizer, all on the sm_100a kernels:

  forward   pack_input -> [conv3x3+bias+ReLU]x2 -> maxpool ... -> convT (GEMM + pixel shuffle) ->
            conv3x3 over (skip || up) with the concat fused into the K loop ... -> head (1x1 conv + sigmoid +
            BCE/Dice partial sums)
  backward  head backward (+ReLU mask) -> per conv: weight gradient (tcgen05, split-K, deterministic),
            bias gradient, data gradient with the ReLU mask of the producer fused in the epilogue and the
            concat backward fused in the store (two destinations) -> fused max-pool backward + skip add + mask
  update    flat fused Adam / RMSprop (or the fused all-reduce + optimizer kernel when a team is attached),
            then one kernel that refreshes the bf16 kernel-side weight layouts

Parameters live in ONE flat fp32 arena laid out in forward execution order, so gradients become ready from
the back of the arena to the front during backward and gradient buckets are contiguous ranges.  The
nn.Module's parameters are views into that arena: state_dict(), checkpoints and the torch-op path all see
the same storage.

The network is nine blocks (enc1..enc4, mid, dec1..dec4+head).  An engine can own any contiguous range of
them (`blocks=(lo, hi)`, pipeline stages) and several engines can share one parameter core (`shared=`, one
engine per in-flight micro-batch, gradients accumulated with `accumulate=True`).
"""
from __future__ import annotations

import ctypes as C
import os

from dataclasses import dataclass
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..models.unet import UNet
from ..ops import native as N
from .program import Program

NUM_BLOCKS = 9


@dataclass
class ParamSlot:
    name: str
    offset: int
    numel: int
    shape: Tuple[int, ...]


def block_param_prefixes(block: int) -> List[str]:
    if block < 4:
        return [f"encoder.conv{block + 1}."]
    if block == 4:
        return ["mid."]
    i = block - 4
    out = [f"decoder.deconv{i}.", f"decoder.conv{i}."]
    if block == NUM_BLOCKS - 1:
        out.append("segmap.")
    return out


def _forward_order_names(model: UNet, blocks: Tuple[int, int] = (0, NUM_BLOCKS)) -> List[str]:
    names = []
    for i in range(1, 5):
        for j in (0, 2):
            names += [f"encoder.conv{i}.conv_block.{j}.weight", f"encoder.conv{i}.conv_block.{j}.bias"]
    for j in (0, 2):
        names += [f"mid.conv_block.{j}.weight", f"mid.conv_block.{j}.bias"]
    for i in range(1, 5):
        names += [f"decoder.deconv{i}.weight", f"decoder.deconv{i}.bias"]
        for j in (0, 2):
            names += [f"decoder.conv{i}.conv_block.{j}.weight", f"decoder.conv{i}.conv_block.{j}.bias"]
    names += ["segmap.weight", "segmap.bias"]
    have = dict(model.named_parameters())
    assert set(names) == set(have.keys()), "unexpected parameter set for the reference UNet"
    prefixes = [p for b in range(blocks[0], blocks[1]) for p in block_param_prefixes(b)]
    return [n for n in names if any(n.startswith(p) for p in prefixes)]


_NVTX = os.environ.get("DPT_NVTX", "0") == "1"
# program events that tell the input-staging stream when the step has finished reading its input buffers
FUSE_POOL = os.environ.get("DPT_FUSE_POOL", "1") != "0"
EV_IMAGES_FREE, EV_TARGET_FREE = 5, 6
# Backward on two streams: the data-gradient chain (the critical path) stays on the step's stream, every weight gradient
# (tcgen05 kernel + split-K reduction) goes to a second stream as soon as its output gradient exists.  Both kinds of
# kernels are one-CTA-per-SM grids with a static tile schedule, so alone each leaves SMs idle whenever its tile count is not
# a multiple of the SM count (mid.2: 300 tiles on 148 SMs = three rounds for two rounds of work; its weight gradient has
# 128 CTAs): with two independent kernels in flight the other kernel's CTAs take those SMs.
EV_DZ_READY, EV_WGRADS_DONE = 3, 4
WGRAD_STREAM = os.environ.get("DPT_WGRAD_STREAM", "1") != "0"
# Early update (DPT_EARLY_UPDATE=1; single GPU, with the two-stream backward): once the bottleneck's data gradient has run,
# nothing in the step reads the decoder / bottleneck / head weights any more and their gradients are complete, so their
# optimizer update and bf16 re-layout (85 % of the parameters) can be enqueued on the second stream under the encoder's
# backward instead of after it.  Same arithmetic per element, bit-identical results (tests/test_gpu_engine.py) - and
# measured SLOWER: 930.6 -> 924.2 images/s.  The update kernels are HBM-bound grids of small CTAs; while they sit on the
# SMs the one-CTA-per-SM data-gradient kernels of the critical chain cannot become resident.  Off by default.
EARLY_UPDATE = os.environ.get("DPT_EARLY_UPDATE", "0") == "1"


class ParamArena:
    """Flat fp32 parameter / gradient arenas with the module's parameters (of the owned blocks) as views."""

    def __init__(self, model: UNet, device: torch.device, params: Optional[torch.Tensor] = None,
                 grads: Optional[torch.Tensor] = None, blocks: Tuple[int, int] = (0, NUM_BLOCKS)):
        names = _forward_order_names(model, blocks)
        have = dict(model.named_parameters())
        self.slots: Dict[str, ParamSlot] = {}
        off = 0
        for n in names:
            p = have[n]
            self.slots[n] = ParamSlot(n, off, p.numel(), tuple(p.shape))
            off += (p.numel() + 3) // 4 * 4  # 16-byte aligned slots
        self.numel = off
        self.device = device
        self.params = params if params is not None else torch.zeros(off, dtype=torch.float32, device=device)
        self.grads = grads if grads is not None else torch.zeros(off, dtype=torch.float32, device=device)
        assert self.params.numel() >= off and self.grads.numel() >= off
        with torch.no_grad():
            for n in names:
                s = self.slots[n]
                view = self.params[s.offset:s.offset + s.numel].view(s.shape)
                view.copy_(have[n].detach().to(device=device, dtype=torch.float32))
                have[n].data = view
                have[n].grad = self.grads[s.offset:s.offset + s.numel].view(s.shape)

    def p(self, name: str) -> torch.Tensor:
        s = self.slots[name]
        return self.params[s.offset:s.offset + s.numel].view(s.shape)

    def g(self, name: str) -> torch.Tensor:
        s = self.slots[name]
        return self.grads[s.offset:s.offset + s.numel].view(s.shape)


class NativeUNetEngine:
    """Forward / train-step executor for `UNet(variant='ref')` at a fixed (batch, height, width)."""

    WIDTHS = (32, 64, 128, 256)

    def __init__(self, model: UNet, batch: int, height: int, width: int, device, *, lr: float = 1e-4,
                 weight_decay: float = 1e-8, optimizer: str = "adam", betas=(0.9, 0.999), eps: float = 1e-8,
                 rms_alpha: float = 0.99, rms_momentum: float = 0.0, loss_scale: Optional[float] = None,
                 dice_weight: float = 1.0, dice_scale: float = 1.0, target_dtype: str = "float32",
                 input_dtype: str = "float32",
                 params: Optional[torch.Tensor] = None, grads: Optional[torch.Tensor] = None, build_update: bool = True,
                 total_pixels: Optional[int] = None, io_tensors: Optional[Dict[str, torch.Tensor]] = None,
                 pre_finalize=None, blocks: Tuple[int, int] = (0, NUM_BLOCKS), shared: "Optional[NativeUNetEngine]" = None,
                 accumulate: bool = False, finalize: bool = True, grads_ready_hook=None):
        if model.variant != "ref":
            raise ValueError("NativeUNetEngine drives the reference variant (variant='milesial' runs on the torch-op path)")
        if height < 16 or width < 16:
            raise ValueError("height and width must be at least 16 (four 2x2 poolings)")
        self.model = model
        self.B, self.H, self.W = batch, height, width
        # Encoder level l works at floor(H / 2^l); the decoder doubles back up from level 4, so for H, W that are not
        # multiples of 16 its tensors are smaller than the skip tensors, which are center-cropped (model/unet_parts.py:
        # 58-74) - here as tensor-map windows - and the output is 16*floor(H/16) x 16*floor(W/16).  The reference cannot
        # TRAIN at such a size (its loss compares that output with the full-size mask and raises); neither does this
        # engine: it builds the forward (inference) program only.
        self.enc_hw = [(height >> l, width >> l) for l in range(5)]
        self.dec_hw = [((height >> 4) << (4 - l), (width >> 4) << (4 - l)) for l in range(5)]
        self.inference_only = self.enc_hw != self.dec_hw
        self.device = torch.device(device)
        self.lr, self.weight_decay, self.eps = lr, weight_decay, eps
        self.optimizer = optimizer
        self.betas = betas
        self.rms_alpha, self.rms_momentum = rms_alpha, rms_momentum
        self.loss_scale = float(batch if loss_scale is None else loss_scale)  # reference: (batch_size * loss).backward()
        # the reference treats dice_weight as a switch (utils/utils.py:16); dice_scale weights a micro-batch's term
        self.dice_weight = (1.0 if dice_weight else 0.0) * dice_scale
        self.target_dtype = 1 if target_dtype == "uint8" else 0
        # "uint8": batches arrive as raw pixel values and are normalised by the packing kernel (1/4 of the H2D bytes)
        self.input_uint8 = input_dtype == "uint8"
        self.wants_float_masks = self.target_dtype == 0
        self.total_pixels = int(total_pixels if total_pixels is not None else batch * height * width)
        self.io_tensors = io_tensors or {}  # externally owned images / target / prob / head_partial / loss_state
        self.pre_finalize = pre_finalize
        # called as hook(program, layer_name) right after the ops that complete a layer's parameter gradients have been
        # appended to the backward program (DDP uses it to launch bucket collectives on the aux stream)
        self.grads_ready_hook = grads_ready_hook
        self.blocks = (int(blocks[0]), int(blocks[1]))
        self.accumulate = accumulate
        self.finalize = finalize
        self.has_head = self.blocks[1] == NUM_BLOCKS
        self.has_input = self.blocks[0] == 0
        # whole-model, one-batch-per-step engines only: a pipeline stage interleaves micro-batches on its buffers and DDP's
        # overlap mode owns the second stream
        self.wgrad_concurrent = (WGRAD_STREAM and grads_ready_hook is None and shared is None and not accumulate
                                 and self.blocks == (0, NUM_BLOCKS))
        # the incomplete last batch of an epoch (utils/train_utils.py:40 trains on it) runs on a second engine of that batch
        # size that shares this one's parameters, built on first use: whole-model engines only
        self.supports_partial_batch = shared is None and self.blocks == (0, NUM_BLOCKS) and not io_tensors
        self._partial: Dict[int, "NativeUNetEngine"] = {}
        self._partial_kwargs = dict(lr=lr, weight_decay=weight_decay, optimizer=optimizer, betas=betas, eps=eps,
                                    rms_alpha=rms_alpha, rms_momentum=rms_momentum, loss_scale=self.loss_scale,
                                    dice_weight=dice_weight, dice_scale=dice_scale, target_dtype=target_dtype,
                                    input_dtype=input_dtype)
        N.lib()
        with torch.cuda.device(self.device):
            if shared is None:
                self.arena = ParamArena(model, self.device, params, grads, self.blocks)
                self._alloc_core()
                self._build_weights()
            else:  # another micro-batch of the same stage: same parameters, weights, optimizer state, scratch
                for k in ("arena", "opt_state", "m", "v", "csum_scratch", "wf", "wd", "bias4", "prep_table", "prep_entries"):
                    setattr(self, k, getattr(shared, k))
            self._alloc_buffers()
            self.wgrad_scratch = shared.wgrad_scratch if shared is not None else torch.empty(
                max(16, self._wgrad_scratch_needed()), dtype=torch.uint8, device=self.device)
            self.fwd = Program(self.device)
            if self.wgrad_concurrent and not self.inference_only:
                self.aux_stream = torch.cuda.Stream(device=self.device)
            if not self.inference_only:
                self._build_forward(self.fwd, with_loss=True)
                self.fwd_end = len(self.fwd)
                self._build_backward(self.fwd)
                self.bwd_end = len(self.fwd)
                if build_update and shared is None:
                    self._build_update(self.fwd)
            else:
                self.fwd_end = self.bwd_end = 0
                self.supports_partial_batch = False
            self.step_end = len(self.fwd)
            # what train_step() runs: the same ops with the early update woven into the backward when that applies
            # (self.fwd keeps the plain forward | backward | update ranges for every other entry point)
            self.step_prog = self.fwd
            self.early_update = (EARLY_UPDATE and self.wgrad_concurrent and build_update and not self.inference_only
                                 and self.arena.numel > 0)
            if self.early_update:
                self.step_prog = Program(self.device)
                self._build_forward(self.step_prog, with_loss=True, set_ranges=False)
                self._build_backward(self.step_prog, early=True)
                self._build_update(self.step_prog, early=True)
            self.infer = Program(self.device)
            self._build_forward(self.infer, with_loss=False)
            if shared is None:
                self.refresh_weights()

    def owns(self, block: int) -> bool:
        return self.blocks[0] <= block < self.blocks[1]

    # -------------------------------------------------------------------------------------------------
    def _buf(self, lvl: int, c: int, dec: bool = False) -> torch.Tensor:
        h, w = (self.dec_hw if dec else self.enc_hw)[lvl]
        return torch.zeros(self.B, h, w, c, dtype=torch.bfloat16, device=self.device)

    def _alloc_core(self) -> None:
        dev = self.device
        self.opt_state = torch.zeros(8, device=dev)
        self.m = torch.zeros(max(4, self.arena.numel), device=dev)
        self.v = torch.zeros(max(4, self.arena.numel), device=dev)
        self.csum_scratch = torch.empty(int(N.lib().dpt_channel_sum_scratch_bytes(512)), dtype=torch.uint8, device=dev)
        N.check(N.lib().dpt_opt_init(self.opt_state.data_ptr(), self.lr, 0, self.betas[0], self.betas[1], N.stream_ptr()))

    def _alloc_buffers(self) -> None:
        """Activations / activation gradients of the owned blocks plus the tensors that cross into them."""
        dev, B, H, W = self.device, self.B, self.H, self.W
        tdt = torch.uint8 if self.target_dtype == 1 else torch.float32
        io = self.io_tensors
        self.images = io["images"] if "images" in io else torch.zeros(
            B, 3, H, W, dtype=torch.uint8 if self.input_uint8 else torch.float32, device=dev)
        self.target = io["target"] if "target" in io else torch.zeros(B, H, W, dtype=tdt, device=dev)
        self.prob = io["prob"] if "prob" in io else torch.zeros(B, 1, *self.dec_hw[0], dtype=torch.float32, device=dev)
        assert self.images.is_contiguous() and self.target.is_contiguous() and self.prob.is_contiguous()
        a: Dict[str, torch.Tensor] = {}
        g: Dict[str, torch.Tensor] = {}
        lo, hi = self.blocks

        def need_act(block_producer: int, block_consumer: int) -> bool:
            return self.owns(block_producer) or self.owns(block_consumer)

        if self.has_input:
            a["x0"] = self._buf(0, 16)
        for l, c in enumerate(self.WIDTHS):
            if self.owns(l):
                a[f"e{l}a"], g[f"e{l}a"] = self._buf(l, c), self._buf(l, c)
            if need_act(l, NUM_BLOCKS - 1 - l):  # skip tensor e{l}b: produced by block l, consumed by block 8-l
                a[f"e{l}b"], g[f"skip{l}"] = self._buf(l, c), self._buf(l, c)
            if need_act(l, l + 1):  # pooled tensor p{l}: main flow l -> l+1
                a[f"p{l}"], g[f"p{l}"] = self._buf(l + 1, c), self._buf(l + 1, c)
        if self.owns(4):
            a["ma"], g["ma"] = self._buf(4, 512), self._buf(4, 512)
        if need_act(4, 5):
            a["mb"], g["mb"] = self._buf(4, 512), self._buf(4, 512)
        for i in range(1, 5):
            lvl, c = 4 - i, self.WIDTHS[4 - i]
            if self.owns(4 + i):
                for k in (f"u{i}", f"d{i}a"):
                    a[k], g[k] = self._buf(lvl, c, True), self._buf(lvl, c, True)
            if need_act(4 + i, 5 + i) if i < 4 else self.owns(8):
                a[f"d{i}b"], g[f"d{i}b"] = self._buf(lvl, c, True), self._buf(lvl, c, True)
        if self.inference_only:
            g = {}  # no backward at this size: the gradient buffers are dropped again
        self.act, self.grad = a, g
        hb = N.lib().dpt_head_blocks()
        self.head_blocks = hb
        self.head_partial = io["head_partial"] if "head_partial" in io else torch.zeros(hb, 4, device=dev)
        self.head_partial_w = torch.zeros(hb, 33, device=dev)
        self.loss_state = io["loss_state"] if "loss_state" in io else torch.zeros(8, device=dev)

    # flow name (parallel/pipe.py) -> buffer key
    @staticmethod
    def flow_key(flow: str) -> str:
        kind, idx = flow[:4], int(flow[4:])
        if kind == "skip":
            return f"e{idx}b"
        if idx < 4:
            return f"p{idx}"
        return "mb" if idx == 4 else f"d{idx - 4}b"

    @staticmethod
    def flow_grad_key(flow: str) -> str:
        kind, idx = flow[:4], int(flow[4:])
        if kind == "skip":
            return f"skip{idx}"
        return NativeUNetEngine.flow_key(flow)

    # (module prefix, cin, cout, kind, block)  kind 0 = conv3x3, 1 = convT
    def _layers(self) -> List[Tuple[str, int, int, int]]:
        L: List[Tuple[str, int, int, int, int]] = []
        chans = [3] + list(self.WIDTHS)
        for i in range(4):
            L += [(f"encoder.conv{i + 1}.conv_block.0", chans[i], chans[i + 1], 0, i),
                  (f"encoder.conv{i + 1}.conv_block.2", chans[i + 1], chans[i + 1], 0, i)]
        L += [("mid.conv_block.0", 256, 512, 0, 4), ("mid.conv_block.2", 512, 512, 0, 4)]
        for i in range(1, 5):
            c = self.WIDTHS[4 - i]
            L += [(f"decoder.deconv{i}", 2 * c, c, 1, 4 + i), (f"decoder.conv{i}.conv_block.0", 2 * c, c, 0, 4 + i),
                  (f"decoder.conv{i}.conv_block.2", c, c, 0, 4 + i)]
        return [(n, ci, co, k) for n, ci, co, k, b in L if self.owns(b)]

    def _build_weights(self) -> None:
        dev = self.device
        self.wf: Dict[str, torch.Tensor] = {}
        self.wd: Dict[str, Optional[torch.Tensor]] = {}
        self.bias4: Dict[str, torch.Tensor] = {}
        entries = []
        for name, cin, cout, kind in self._layers():
            w = self.arena.p(name + ".weight")
            b = self.arena.p(name + ".bias")
            if kind == 0:
                cin_pad = 16 if cin < 16 else cin
                self.wf[name] = torch.zeros(9, cout, cin_pad, dtype=torch.bfloat16, device=dev)
                self.wd[name] = torch.zeros(9, cin, cout, dtype=torch.bfloat16, device=dev) if cin >= 16 else None
                entries.append(N.PrepEntry(w.data_ptr(), None, self.wf[name].data_ptr(), N.ptr(self.wd[name]), None, cout,
                                           cin, cin_pad, 0))
            else:
                self.wf[name] = torch.zeros(1, 4 * cout, cin, dtype=torch.bfloat16, device=dev)
                self.wd[name] = torch.zeros(4, cin, cout, dtype=torch.bfloat16, device=dev)
                self.bias4[name] = torch.zeros(4 * cout, device=dev)
                entries.append(N.PrepEntry(w.data_ptr(), b.data_ptr(), self.wf[name].data_ptr(), self.wd[name].data_ptr(),
                                           self.bias4[name].data_ptr(), cout, cin, cin, 1))
        self.prep_table = N.prep_weights(entries, dev) if entries else torch.zeros(1, dtype=torch.uint8, device=dev)
        self.prep_entries = len(entries)

    def refresh_weights(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        """Re-derive the bf16 kernel-side weights from the fp32 arena (after load_state_dict / optimizer step)."""
        if self.prep_entries:
            N.run_prep_weights(self.prep_table, self.prep_entries, stream)

    # -------------------------------------------------------------------------------------------------
    def _conv(self, prog: Program, name: str, srcs: Sequence[torch.Tensor], dst: torch.Tensor, lvl: int, label: str,
              pooled: Optional[torch.Tensor] = None):
        h, w = int(dst.shape[1]), int(dst.shape[2])
        # a source larger than the output (skip tensor at a size that is not a multiple of 16): CenterCrop as a window,
        # corner at round((Hs - h) / 2) like torchvision.transforms.CenterCrop
        offs = [None if (s.shape[1], s.shape[2]) == (h, w) else (int(round((s.shape[1] - h) / 2.0)), int(round((s.shape[2] - w) / 2.0)))
                for s in srcs]
        plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=list(srcs),
                          weight=self.wf[name], cout=dst.shape[-1], dsts=[dst], bias=self.arena.p(name + ".bias"),
                          relu=True, src_offsets=offs if any(o is not None for o in offs) else None, pooled=pooled)
        prog.conv(plan, label)

    def _build_forward(self, prog: Program, with_loss: bool, set_ranges: bool = True) -> None:
        a = self.act
        # op index range of every owned block inside `prog` (pipeline stages run, wait and send block by block)
        ranges: Dict[int, Tuple[int, int]] = {}
        if set_ranges and with_loss:
            self.fwd_block_range = ranges
        elif set_ranges:
            self.infer_block_range = ranges
        mark = len(prog)

        def close(block: int) -> None:
            nonlocal mark
            ranges[block] = (mark, len(prog))
            mark = len(prog)

        if self.has_input:
            prog.pack_input(self.images, a["x0"])
            if with_loss:
                prog.record(EV_IMAGES_FREE, label="input images consumed")
        for l in range(4):
            if not self.owns(l):
                continue
            x = a["x0"] if l == 0 else a[f"p{l - 1}"]
            base = f"encoder.conv{l + 1}.conv_block"
            self._conv(prog, base + ".0", [x], a[f"e{l}a"], l, f"fwd enc{l + 1}.0")
            # the encoder's MaxPool2d(2, 2) is a second store of the same epilogue when the extent is even (training
            # sizes always are); odd extents (inference at arbitrary sizes) keep the floor-mode pooling kernel
            eh, ew = int(a[f"e{l}b"].shape[1]), int(a[f"e{l}b"].shape[2])
            fuse = FUSE_POOL and eh % 2 == 0 and ew % 2 == 0
            self._conv(prog, base + ".2", [a[f"e{l}a"]], a[f"e{l}b"], l, f"fwd enc{l + 1}.2" + (" + pool" if fuse else ""),
                       pooled=a[f"p{l}"] if fuse else None)
            if not fuse:
                prog.maxpool(a[f"e{l}b"], a[f"p{l}"], f"fwd pool{l + 1}")
            close(l)
        if self.owns(4):
            self._conv(prog, "mid.conv_block.0", [a["p3"]], a["ma"], 4, "fwd mid.0")
            self._conv(prog, "mid.conv_block.2", [a["ma"]], a["mb"], 4, "fwd mid.2")
            close(4)
        for i in range(1, 5):
            if not self.owns(4 + i):
                continue
            lvl = 4 - i
            x = a["mb"] if i == 1 else a[f"d{i - 1}b"]
            name = f"decoder.deconv{i}"
            plan = N.ConvPlan(mode=N.CONV_1X1, out_mode=N.OUT_SHUFFLE2, batch=self.B, h=int(x.shape[1]), w=int(x.shape[2]),
                              srcs=[x], weight=self.wf[name], cout=4 * a[f"u{i}"].shape[-1],
                              dsts=[a[f"u{i}"]], bias=self.bias4[name])
            prog.conv(plan, f"fwd deconv{i}")
            base = f"decoder.conv{i}.conv_block"
            self._conv(prog, base + ".0", [a[f"e{lvl}b"], a[f"u{i}"]], a[f"d{i}a"], lvl, f"fwd dec{i}.0 (skip||up)")
            self._conv(prog, base + ".2", [a[f"d{i}a"]], a[f"d{i}b"], lvl, f"fwd dec{i}.2")
            if i < 4:
                close(4 + i)
        if not self.has_head:
            return
        w, b = self.arena.p("segmap.weight"), self.arena.p("segmap.bias")
        if with_loss:
            prog.head_fwd(a["d4b"], w, b, self.target, self.target_dtype, self.prob, self.head_partial, "fwd head+loss")
            if self.pre_finalize is not None:  # e.g. DataParallel: all-reduce the loss statistics across replicas
                self.pre_finalize(prog)
            if self.finalize:
                prog.head_finalize([self.head_partial], self.head_blocks, self.total_pixels, self.dice_weight,
                                   self.loss_state, None, 0, "loss finalize")
        else:
            prog.head_fwd(a["d4b"], w, b, None, 0, self.prob, self.head_partial, "fwd head")
        close(NUM_BLOCKS - 1)

    # -------------------------------------------------------------------------------------------------
    def _wgrad_scratch_needed(self) -> int:
        need = 0
        B = self.B
        if self.inference_only:
            return 0
        for name, cin, cout, kind in self._layers():
            lvl = self._level_of(name)
            h, w = self.H >> lvl, self.W >> lvl
            if kind == 0:
                xcs = [cin // 2, cin // 2] if name.startswith("decoder.conv") and name.endswith(".0") else [max(cin, 16)]
                need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_3X3, batch=B, h=h, w=w, z_c=cout, x_cs=xcs))
            else:
                need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_2X2S2, batch=B, h=h, w=w, z_c=cin, x_cs=[cout]))
        return need

    def _level_of(self, name: str) -> int:
        """Resolution level of the GEMM pixel space of a layer (convT: its INPUT resolution)."""
        if name.startswith("encoder.conv"):
            return int(name[len("encoder.conv")]) - 1
        if name.startswith("mid"):
            return 4
        if name.startswith("decoder.deconv"):
            return 5 - int(name[len("decoder.deconv")])
        return 4 - int(name[len("decoder.conv")])

    def _wgrad(self, prog: Program, wp: "N.WgradPlan", name: str) -> None:
        """Append one weight-gradient op.  Two-stream backward: the caller has just recorded EV_DZ_READY behind the kernel
        that produced this layer's output gradient; the second stream waits for it and takes the op."""
        if self.wgrad_concurrent:
            prog.wait(EV_DZ_READY, stream=1, label="wgrad stream: output gradient ready")
            prog.wgrad(wp, f"bwd wgrad+bgrad {name}", stream=1)
        else:
            prog.wgrad(wp, f"bwd wgrad+bgrad {name}")
            if self.grads_ready_hook is not None:
                self.grads_ready_hook(self, prog, name)

    def _conv_bwd(self, prog: Program, name: str, xs: Sequence[torch.Tensor], dz: torch.Tensor, lvl: int,
                  dsts: Optional[Sequence[torch.Tensor]], masks: Sequence[Optional[torch.Tensor]], n_valid: int = 0):
        """Backward of one 3x3 conv given dz (gradient w.r.t. its pre-activation output)."""
        h, w = self.H >> lvl, self.W >> lvl
        cin = sum(x.shape[-1] for x in xs)
        cin_real = n_valid or cin
        wp = N.WgradPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, z=dz, xs=list(xs), dw=self.arena.g(name + ".weight"),
                         strides=(cin_real * 9, 9, 1), scratch=self.wgrad_scratch, n_valid=n_valid,
                         accumulate=self.accumulate, dbias=self.arena.g(name + ".bias"))
        # the bias gradient (column sums of dz) comes out of the same two kernels: the wgrad CTAs add up the dz tiles
        # they already hold in shared memory
        dgrad = None
        if dsts is not None:
            dgrad = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=[dz], weight=self.wd[name], cout=cin,
                               dsts=list(dsts), masks=list(masks))
        if self.wgrad_concurrent:
            # dz is complete on the main stream here; the data gradient is enqueued first (it is what the next layer waits
            # for), the weight gradient fills the SMs it leaves idle
            prog.record(EV_DZ_READY, label="output gradient ready")
            if dgrad is not None:
                prog.conv(dgrad, f"bwd dgrad {name}")
            self._wgrad(prog, wp, name)
        else:
            self._wgrad(prog, wp, name)
            if dgrad is not None:
                prog.conv(dgrad, f"bwd dgrad {name}")

    def _early_slice(self) -> Tuple[int, int]:
        """(first arena element, first re-layout entry) of the parameters updated early: bottleneck, decoder, head."""
        return self.arena.slots["mid.conv_block.0.weight"].offset, 8

    def _build_backward(self, prog: Program, early: bool = False) -> None:
        a, g = self.act, self.grad
        ranges: Dict[int, Tuple[int, int]] = {}
        if not early:
            self.bwd_block_range = ranges
        else:
            prog.opt_tick(self.opt_state)  # the step counter / bias corrections, before the first update kernel
        mark = len(prog)

        def close(block: int) -> None:
            nonlocal mark
            ranges[block] = (mark, len(prog))
            mark = len(prog)

        if self.has_head:
            prog.head_bwd(a["d4b"], self.arena.p("segmap.weight"), self.arena.p("segmap.bias"), self.target,
                          self.target_dtype, self.loss_state, g["d4b"], self.head_partial_w, self.arena.g("segmap.weight"),
                          self.arena.g("segmap.bias"), self.loss_scale, self.dice_weight, True, "bwd head", self.accumulate)
            prog.record(EV_TARGET_FREE, label="target consumed")
            if self.grads_ready_hook is not None:
                self.grads_ready_hook(self, prog, "segmap")
        for i in range(4, 0, -1):
            if not self.owns(4 + i):
                continue
            lvl = 4 - i
            base = f"decoder.conv{i}.conv_block"
            self._conv_bwd(prog, base + ".2", [a[f"d{i}a"]], g[f"d{i}b"], lvl, [g[f"d{i}a"]], [a[f"d{i}a"]])
            self._conv_bwd(prog, base + ".0", [a[f"e{lvl}b"], a[f"u{i}"]], g[f"d{i}a"], lvl,
                           [g[f"skip{lvl}"], g[f"u{i}"]], [None, None])
            # ConvTranspose2d i: input prev (post-ReLU), output gradient g[u_i]
            name = f"decoder.deconv{i}"
            prev, gprev = (a["mb"], g["mb"]) if i == 1 else (a[f"d{i - 1}b"], g[f"d{i - 1}b"])
            hh, ww = self.H >> (lvl + 1), self.W >> (lvl + 1)
            cin, cout = prev.shape[-1], a[f"u{i}"].shape[-1]
            wp = N.WgradPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, z=prev, xs=[g[f"u{i}"]],
                             dw=self.arena.g(name + ".weight"), strides=(cout * 4, 4, 1), scratch=self.wgrad_scratch,
                             accumulate=self.accumulate, dbias=self.arena.g(name + ".bias"))
            plan = N.ConvPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, srcs=[g[f"u{i}"]], weight=self.wd[name],
                              cout=cin, dsts=[gprev], masks=[prev])
            if self.wgrad_concurrent:
                prog.record(EV_DZ_READY, label="output gradient ready")
                prog.conv(plan, f"bwd dgrad {name}")
                self._wgrad(prog, wp, name)
            else:
                self._wgrad(prog, wp, name)
                prog.conv(plan, f"bwd dgrad {name}")
            close(4 + i)
        if self.owns(4):
            self._conv_bwd(prog, "mid.conv_block.2", [a["ma"]], g["mb"], 4, [g["ma"]], [a["ma"]])
            self._conv_bwd(prog, "mid.conv_block.0", [a["p3"]], g["ma"], 4, [g["p3"]], [None])
            close(4)
        for l in range(3, -1, -1):
            if not self.owns(l):
                continue
            base = f"encoder.conv{l + 1}.conv_block"
            prog.pool_bwd(a[f"e{l}b"], g[f"p{l}"], g[f"skip{l}"], g[f"skip{l}"], True, f"bwd pool{l + 1}+skip+relu")
            self._conv_bwd(prog, base + ".2", [a[f"e{l}a"]], g[f"skip{l}"], l, [g[f"e{l}a"]], [a[f"e{l}a"]])
            if early and l == 3:
                # the second stream has just waited for an event recorded behind the bottleneck's data gradient: every
                # reader of the decoder / bottleneck / head weights has finished and their gradients are complete
                self._update_ops(prog, *self._early_slice(), None, None, stream=1, tag=" (bottleneck + decoder + head, early)")
            if l > 0:
                self._conv_bwd(prog, base + ".0", [a[f"p{l - 1}"]], g[f"e{l}a"], l, [g[f"p{l - 1}"]], [None])
            else:
                self._conv_bwd(prog, base + ".0", [a["x0"]], g["e0a"], 0, None, [], n_valid=3)
            if l == 0 and self.wgrad_concurrent:  # join: every weight gradient is complete before the update reads them
                prog.record(EV_WGRADS_DONE, stream=1, label="weight gradients done")
                prog.wait(EV_WGRADS_DONE, label="join the wgrad stream")
            close(l)

    def _update_ops(self, prog: Program, lo: int, e_lo: int, hi: Optional[int], e_hi: Optional[int], stream: int = 0,
                    tag: str = "") -> None:
        """Optimizer + bf16 re-layout of the arena elements [lo, hi) / re-layout entries [e_lo, e_hi)."""
        hi = self.arena.numel if hi is None else hi
        e_hi = self.prep_entries if e_hi is None else e_hi
        n = hi - lo
        P, G, M = self.arena.params[lo:hi], self.arena.grads[lo:hi], self.m[lo:hi]
        if self.optimizer == "adam":
            prog.adam(P, G, M, self.v[lo:hi], n, self.opt_state, self.eps, self.weight_decay, label="adam" + tag, stream=stream)
        elif self.optimizer == "rmsprop":
            prog.rmsprop(P, G, M, self.v[lo:hi] if self.rms_momentum > 0 else None, n, self.opt_state, self.rms_alpha, self.eps,
                         self.weight_decay, self.rms_momentum, label="rmsprop" + tag, stream=stream)
        else:
            raise ValueError(f"unknown optimizer {self.optimizer!r}")
        esz = C.sizeof(N.PrepEntry)
        prog.prep_weights(self.prep_table[e_lo * esz:], e_hi - e_lo, label="prep_weights" + tag, stream=stream)

    def _build_update(self, prog: Program, early: bool = False) -> None:
        n = self.arena.numel
        if n == 0:
            return
        if early:  # the encoder's parameters; the rest was updated under the encoder's backward (_build_backward)
            lo, e_lo = self._early_slice()
            self._update_ops(prog, 0, 0, lo, e_lo, tag=" (encoder)")
            return
        prog.opt_tick(self.opt_state)
        self._update_ops(prog, 0, 0, None, None)

    # -------------------------------------------------------------------------------------------------
    @property
    def launches_per_step(self) -> int:
        """Kernels launched by one train_step (wgrad / bias-grad / head-backward ops launch two kernels each)."""
        from .program import OP_EVENT_RECORD, OP_STREAM_WAIT

        two = sum(1 for lbl in self.fwd.labels[:self.step_end] if lbl.startswith(("bwd wgrad", "bwd bgrad", "bwd head")))
        events = sum(1 for k in self.fwd.kinds[:self.step_end] if k in (OP_EVENT_RECORD, OP_STREAM_WAIT))
        return self.step_end + two - events

    def set_lr(self, lr: float) -> None:
        self.lr = lr
        N.check(N.lib().dpt_opt_set_lr(self.opt_state.data_ptr(), lr, N.stream_ptr()))

    def load_batch(self, images: Optional[torch.Tensor], masks: Optional[torch.Tensor]) -> None:
        """Copy a batch (device or pinned-host tensors) into the static input buffers, asynchronously."""
        if images is not None:
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
        if masks is not None:
            self.target.copy_(masks.reshape(self.target.shape), non_blocking=True)
        self._input_dirty = True

    # ---- input staging: host -> device copies on their own stream, overlapped with compute ----
    def prefetch(self, images: torch.Tensor, masks: torch.Tensor) -> None:
        """Start copying the NEXT batch (pinned host or device tensors) straight into the engine's input buffers on a side
        stream while the current step computes; the following `train_step()` consumes it.  No staging buffer and no
        device-to-device hop: the copy stream waits for the two program events that say "the step in flight has read its
        images" (right after the packing kernel) and "... its target" (after the head's backward kernel)."""
        if not hasattr(self, "_copy_stream"):
            with torch.cuda.device(self.device):
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._staged = torch.cuda.Event()
        cs = self._copy_stream
        if getattr(self, "_input_dirty", False):  # a load_batch on the compute stream that no step has consumed yet
            cs.wait_stream(torch.cuda.current_stream(self.device))
            self._input_dirty = False
        with torch.cuda.stream(cs):
            self.step_prog.stream_wait(EV_IMAGES_FREE, cs)
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
            self.step_prog.stream_wait(EV_TARGET_FREE, cs)
            self.target.copy_(masks.reshape(self.target.shape), non_blocking=True)
            self._staged.record(cs)
        self._has_staged = True

    def _consume_staged(self) -> None:
        torch.cuda.current_stream(self.device).wait_event(self._staged)
        self._has_staged = False

    def _train_step_partial(self, images: torch.Tensor, masks: torch.Tensor, stream) -> torch.Tensor:
        """A batch smaller than B: forward / loss / backward on a lazily built engine of that size (same parameters,
        weights, gradients arena and scratch; loss scaled by the nominal batch size like the reference's
        `(batch_size * loss).backward()`), then this engine's update ops (optimizer or fused all-reduce + optimizer)."""
        b = int(images.shape[0])
        if not self.supports_partial_batch or not 0 < b < self.B:
            raise ValueError(f"batch of {b} images on an engine built for {self.B}")
        eng = self._partial.get(b)
        if eng is None:
            with torch.cuda.device(self.device):
                eng = NativeUNetEngine(self.model, b, self.H, self.W, self.device, shared=self, build_update=False,
                                       **self._partial_kwargs)
            self._partial[b] = eng
        eng.load_batch(images, masks)
        eng.fwd.run(0, eng.bwd_end, stream)
        self.fwd.run(self.bwd_end, self.step_end, stream, getattr(self, "aux_stream", None))
        return eng.loss_state[4]

    def train_step(self, images: Optional[torch.Tensor] = None, masks: Optional[torch.Tensor] = None,
                   stream: Optional[torch.cuda.Stream] = None) -> torch.Tensor:
        """forward + loss + backward + optimizer.  Returns the loss as a 0-d DEVICE tensor (no sync)."""
        if self.inference_only:
            raise ValueError(f"{self.H}x{self.W} is not a multiple of 16: the output is {self.dec_hw[0]} and the loss against the "
                             "full-size mask is undefined (the reference raises here too); this engine runs forward() only")
        if images is not None and images.shape[0] != self.B:
            return self._train_step_partial(images, masks, stream)
        if images is not None:
            self.load_batch(images, masks)
        elif getattr(self, "_has_staged", False):
            self._consume_staged()
        if _NVTX:  # DPT_NVTX=1: one range per phase for timeline tools (three executor calls instead of one)
            nvtx = torch.cuda.nvtx
            for name, lo, hi in (("dpt forward+loss", 0, self.fwd_end), ("dpt backward", self.fwd_end, self.bwd_end),
                                 ("dpt update", self.bwd_end, self.step_end)):
                nvtx.range_push(name)
                self.fwd.run(lo, hi, stream, getattr(self, "aux_stream", None))
                nvtx.range_pop()
            return self.loss_state[4]
        self.step_prog.run(0, -1 if self.step_prog is not self.fwd else self.step_end, stream, getattr(self, "aux_stream", None))
        self._input_dirty = False
        return self.loss_state[4]

    def forward_backward(self, images=None, masks=None, stream=None) -> torch.Tensor:
        """forward + loss + backward only (gradients left in the arena); used by DDP / DP / tests."""
        if images is not None:
            self.load_batch(images, masks)
        self.fwd.run(0, self.bwd_end, stream)
        return self.loss_state[4]

    def run_forward(self, stream=None) -> None:
        self.fwd.run(0, self.fwd_end, stream)

    def run_backward(self, stream=None) -> None:
        self.fwd.run(self.fwd_end, self.bwd_end, stream)

    def run_update(self, stream=None) -> None:
        self.fwd.run(self.bwd_end, self.step_end, stream)

    def evaluate_batch(self, images: torch.Tensor, masks: torch.Tensor, stream=None) -> torch.Tensor:
        if self.inference_only:
            raise ValueError(f"{self.H}x{self.W} is not a multiple of 16: the loss against the full-size mask is undefined "
                             "(output is {0}x{1}); use forward()".format(*self.dec_hw[0]))
        self.load_batch(images, masks)
        self.fwd.run(0, self.fwd_end, stream)
        return self.loss_state[4]

    def forward(self, images: torch.Tensor, stream=None) -> torch.Tensor:
        """Probabilities [B,1,H,W] fp32 (what the reference's UNet.forward returns)."""
        self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
        self.infer.run(0, -1, stream)
        return self.prob

    def optimizer_state(self) -> dict:
        return {"m": self.m.detach().cpu(), "v": self.v.detach().cpu(), "opt_state": self.opt_state.detach().cpu()}

    def load_optimizer_state(self, sd: dict) -> None:
        self.m.copy_(sd["m"])
        self.v.copy_(sd["v"])
        self.opt_state.copy_(sd["opt_state"])
