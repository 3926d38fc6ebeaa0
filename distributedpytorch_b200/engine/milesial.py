"""Native engine for the north-star model family: `UNet(variant='milesial', n_classes>=2, bilinear=False|True)`
(DoubleConv / Down / Up / OutConv with BatchNorm, model/modelsummary.txt:153-247 of the reference).

Same construction as engine/native_unet.py - one flat fp32 parameter arena with the module's parameters as views, a
native launch program for forward / loss / hand-written backward / optimizer - with the extra pieces of this family:

  conv (tcgen05, no bias) whose epilogue also produces the BatchNorm batch statistics from the fp32 accumulators
      (per-CTA partial sums, no atomics) -> finalize (mean / invstd / running statistics) -> ONE pass that normalises,
      applies ReLU and, for the tensors a Down block consumes, also writes the 2x2 max-pooled tensor
  ConvTranspose (GEMM + pixel shuffle) or bilinear x2;  concat fused into the conv K loop
  OutConv + softmax + cross-entropy + Dice in one pass;  RMSprop (weight_decay 1e-8, momentum 0.999) or Adam
  backward: BatchNorm backward recomputes the ReLU mask from the saved pre-norm activation, so the data-gradient
      convolutions run without a mask; the pool backward adds the skip gradient
  evaluation: BatchNorm folded away - gamma*rsqrt(var+eps) into the bf16 weights, the shift as the conv's epilogue bias

The network is nine blocks (inc, down1..down4, up1..up4+outc), the same cut points as the reference variant's
enc1..enc4, mid, dec1..dec4.  An engine can own any contiguous range of them (`blocks=(lo, hi)`, pipeline stages) and
several engines can share one parameter core (`shared=`, one engine per in-flight micro-batch, gradients accumulated
with `accumulate=True`).  BatchNorm semantics per mode: DataParallel replicas and pipeline micro-batches each normalise
with the statistics of THEIR chunk (what torch's DataParallel / a micro-batched nn.Module does); the running statistics
are the module's buffers (device tensors), updated in place by every forward.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..models.unet import UNet
from ..ops import native as N
from .native_unet import EV_DZ_READY, EV_IMAGES_FREE, EV_TARGET_FREE, EV_WGRADS_DONE, WGRAD_STREAM
from .program import Program

NUM_BLOCKS = 9


class _Unit:
    """One Conv3x3(bias=False) -> BatchNorm -> ReLU unit."""

    def __init__(self, conv_name: str, bn_name: str, cin: int, cout: int, lvl: int, block: int):
        self.conv, self.bn, self.cin, self.cout, self.lvl, self.block = conv_name, bn_name, cin, cout, lvl, block


class NativeMilesialEngine:
    def __init__(self, model: UNet, batch: int, height: int, width: int, device, *, lr: float = 1e-5,
                 weight_decay: float = 1e-8, optimizer: str = "rmsprop", momentum: float = 0.999, alpha: float = 0.99,
                 eps: float = 1e-8, betas=(0.9, 0.999), loss_scale: float = 1.0, params: Optional[torch.Tensor] = None,
                 grads: Optional[torch.Tensor] = None, build_update: bool = True, total_pixels: Optional[int] = None,
                 io_tensors: Optional[Dict[str, torch.Tensor]] = None, pre_finalize=None,
                 blocks: Tuple[int, int] = (0, NUM_BLOCKS), shared: "Optional[NativeMilesialEngine]" = None,
                 accumulate: bool = False, finalize: bool = True, grads_ready_hook=None):
        if model.variant != "milesial" or model.n_classes < 2:
            raise ValueError("NativeMilesialEngine needs UNet(variant='milesial', n_classes>=2)")
        if height % 16 or width % 16:
            raise ValueError("height and width must be multiples of 16")
        self.model = model
        self.K = model.n_classes
        self.bilinear = model.bilinear
        self.B, self.H, self.W = batch, height, width
        self.device = torch.device(device)
        self.lr, self.weight_decay, self.eps = lr, weight_decay, eps
        self.optimizer, self.momentum, self.alpha, self.betas = optimizer, momentum, alpha, betas
        self.loss_scale = loss_scale
        self.target_dtype = 1  # class ids as uint8
        self.input_uint8 = False
        self.wants_float_masks = False
        self.total_pixels = int(total_pixels if total_pixels is not None else batch * height * width)
        self.io_tensors = io_tensors or {}
        self.pre_finalize = pre_finalize
        self.grads_ready_hook = grads_ready_hook
        self.blocks = (int(blocks[0]), int(blocks[1]))
        self.accumulate = accumulate
        self.finalize = finalize
        self.has_head = self.blocks[1] == NUM_BLOCKS
        self.has_input = self.blocks[0] == 0
        # two-stream backward (see native_unet.py): weight gradients on a second stream, whole-model engines only
        self.wgrad_concurrent = (WGRAD_STREAM and grads_ready_hook is None and shared is None and not accumulate
                                 and self.blocks == (0, NUM_BLOCKS))
        # incomplete last batch of an epoch: a second engine of that size sharing this one's parameters (see NativeUNetEngine)
        self.supports_partial_batch = shared is None and self.blocks == (0, NUM_BLOCKS) and not io_tensors
        self._partial: Dict[int, "NativeMilesialEngine"] = {}
        self._partial_kwargs = dict(lr=lr, weight_decay=weight_decay, optimizer=optimizer, momentum=momentum, alpha=alpha,
                                    eps=eps, betas=betas, loss_scale=loss_scale)
        f = 2 if self.bilinear else 1
        self.enc_c = [64, 128, 256, 512, 1024 // f]          # inc, down1..down4 output channels
        self.dec_out = [512 // f, 256 // f, 128 // f, 64]     # up1..up4 output channels
        N.lib()
        with torch.cuda.device(self.device):
            if shared is None:
                for b in model.buffers():
                    b.data = b.data.to(self.device)
                self._layout_params(params, grads)
                self._alloc_core()
                self._build_weights()
            else:
                for k in ("slots", "numel", "params", "grads", "opt_state", "s0", "s1", "bn_scratch", "csum_scratch", "wf",
                          "wd", "bias4", "prep_table", "prep_entries", "eval_table", "eval_bn", "core_state"):
                    setattr(self, k, getattr(shared, k))
            self._alloc()
            self.wgrad_scratch = shared.wgrad_scratch if shared is not None else torch.empty(
                max(16, self._wgrad_scratch_needed()), dtype=torch.uint8, device=self.device)
            self.fwd = Program(self.device)
            if self.wgrad_concurrent:
                self.aux_stream = torch.cuda.Stream(device=self.device)
            self._build_forward(self.fwd, train=True)
            self.fwd_end = len(self.fwd)
            self._build_backward(self.fwd)
            self.bwd_end = len(self.fwd)
            if build_update and shared is None:
                self._build_update(self.fwd)
            self.step_end = len(self.fwd)
            self.evalp = Program(self.device)
            self._build_forward(self.evalp, train=False)
            if shared is None:
                self.eval_prep = Program(self.device)
                self._build_eval_prep(self.eval_prep)
                self.refresh_weights()

    def owns(self, block: int) -> bool:
        return self.blocks[0] <= block < self.blocks[1]

    # ---- structure ----------------------------------------------------------------------------------------
    def _units(self) -> List[_Unit]:
        U: List[_Unit] = []
        c = self.enc_c
        U += [_Unit("inc.double_conv.0", "inc.double_conv.1", 3, 64, 0, 0), _Unit("inc.double_conv.3", "inc.double_conv.4", 64, 64, 0, 0)]
        for l in range(1, 5):
            base = f"down{l}.maxpool_conv.1.double_conv"
            U += [_Unit(base + ".0", base + ".1", c[l - 1], c[l], l, l), _Unit(base + ".3", base + ".4", c[l], c[l], l, l)]
        for i in range(1, 5):
            lvl = 4 - i
            cin = c[lvl] + self.up_ch(i)
            cout = self.dec_out[i - 1]
            mid = cout if not self.bilinear else cin // 2
            base = f"up{i}.conv.double_conv"
            U += [_Unit(base + ".0", base + ".1", cin, mid, lvl, 4 + i), _Unit(base + ".3", base + ".4", mid, cout, lvl, 4 + i)]
        return U

    def _owned_units(self) -> List[_Unit]:
        return [u for u in self._units() if self.owns(u.block)]

    def x_ch(self, i: int) -> int:
        """Channels of the tensor entering Up block i (down4's output, then the previous Up block's output)."""
        return self.enc_c[4] if i == 1 else self.dec_out[i - 2]

    def up_ch(self, i: int) -> int:
        """Channels after the up-sampling of Up block i: ConvTranspose halves them, bilinear keeps them."""
        return self.x_ch(i) if self.bilinear else self.x_ch(i) // 2

    @staticmethod
    def block_param_prefixes(block: int) -> List[str]:
        if block == 0:
            return ["inc."]
        if block <= 4:
            return [f"down{block}."]
        out = [f"up{block - 4}."]
        if block == NUM_BLOCKS - 1:
            out.append("outc.")
        return out

    def _param_order(self) -> List[str]:
        """Every parameter of the model in forward-execution order (gradients become ready back to front)."""
        units = self._units()
        order: List[str] = []
        for u in units[:10]:
            order += [u.conv + ".weight", u.bn + ".weight", u.bn + ".bias"]
        for i in range(1, 5):
            if not self.bilinear:
                order += [f"up{i}.up.weight", f"up{i}.up.bias"]
            for u in units[10 + 2 * (i - 1):12 + 2 * (i - 1)]:
                order += [u.conv + ".weight", u.bn + ".weight", u.bn + ".bias"]
        order += ["outc.conv.weight", "outc.conv.bias"]
        return order

    def _layout_params(self, params, grads) -> None:
        have = dict(self.model.named_parameters())
        order = self._param_order()
        assert set(order) == set(have.keys()), sorted(set(have.keys()) ^ set(order))[:5]
        prefixes = [p for b in range(self.blocks[0], self.blocks[1]) for p in self.block_param_prefixes(b)]
        order = [n for n in order if any(n.startswith(p) for p in prefixes)]
        self.slots: Dict[str, Tuple[int, int, Tuple[int, ...]]] = {}
        off = 0
        for n in order:
            p = have[n]
            self.slots[n] = (off, p.numel(), tuple(p.shape))
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        dev = self.device
        self.params = params if params is not None else torch.zeros(max(off, 4), device=dev)
        self.grads = grads if grads is not None else torch.zeros(max(off, 4), device=dev)
        assert self.params.numel() >= off and self.grads.numel() >= off
        with torch.no_grad():
            for n in order:
                o, k, shp = self.slots[n]
                view = self.params[o:o + k].view(shp)
                view.copy_(have[n].detach().to(dev, torch.float32))
                have[n].data = view
                have[n].grad = self.grads[o:o + k].view(shp)

    def p(self, name):
        o, k, shp = self.slots[name]
        return self.params[o:o + k].view(shp)

    def g(self, name):
        o, k, shp = self.slots[name]
        return self.grads[o:o + k].view(shp)

    def _buffer(self, name: str) -> torch.Tensor:
        return dict(self.model.named_buffers())[name]

    def _buf(self, lvl: int, c: int) -> torch.Tensor:
        return torch.zeros(self.B, self.H >> lvl, self.W >> lvl, c, dtype=torch.bfloat16, device=self.device)

    def _alloc_core(self) -> None:
        """State shared by all micro-batch engines of one stage: optimizer state and kernel scratch."""
        dev = self.device
        L = N.lib()
        self.opt_state = torch.zeros(8, device=dev)
        self.s0 = torch.zeros(max(self.numel, 4), device=dev)
        self.s1 = torch.zeros(max(self.numel, 4), device=dev)
        # BatchNorm scratch: backward partials [296][2][C] + coefficients [2][C]; the forward statistics rows the conv
        # epilogue writes ([grid <= SMs][2][C]) fit in the same buffer
        self.bn_scratch = torch.empty(int(L.dpt_bn_scratch_bytes(1024)) + 2 * 1024 * 4, dtype=torch.uint8, device=dev)
        self.csum_scratch = torch.empty(int(L.dpt_channel_sum_scratch_bytes(1024)), dtype=torch.uint8, device=dev)
        self.core_state = {"weights": "train", "owner": self}
        N.check(L.dpt_opt_init(self.opt_state.data_ptr(), self.lr, 0, self.betas[0], self.betas[1], N.stream_ptr()))

    # flow name (parallel/pipe.py) -> key of self.act / self.grad
    def flow_key(self, flow: str) -> str:
        kind, idx = flow[:4], int(flow[4:])
        U = self._units()
        if kind == "skip":
            return U[2 * idx + 1].conv
        if idx < 4:
            return f"pool{idx + 1}"
        return U[9].conv if idx == 4 else U[11 + 2 * (idx - 5)].conv

    flow_grad_key = flow_key

    def _alloc(self) -> None:
        dev, B, H, W = self.device, self.B, self.H, self.W
        io = self.io_tensors
        self.images = io["images"] if "images" in io else torch.zeros(B, 3, H, W, device=dev)
        self.target = io["target"] if "target" in io else torch.zeros(B, H, W, dtype=torch.uint8, device=dev)
        self.logits = io["logits"] if "logits" in io else torch.zeros(B, self.K, H, W, device=dev)
        assert self.images.is_contiguous() and self.target.is_contiguous() and self.logits.is_contiguous()
        self.input_uint8 = self.images.dtype == torch.uint8
        U = self._units()
        self.raw: Dict[str, torch.Tensor] = {}    # pre-norm conv outputs (owned units)
        self.g_raw: Dict[str, torch.Tensor] = {}  # gradient w.r.t. raw (dz of the conv)
        self.bnbuf: Dict[str, torch.Tensor] = {}
        self.act: Dict[str, torch.Tensor] = {}    # post BN+ReLU activations, pooled tensors ("pool{l}"), up-sampled ("up{i}")
        self.grad: Dict[str, torch.Tensor] = {}   # gradients of the same
        if self.has_input:
            self.x0 = self._buf(0, 16)

        def both(key, lvl, c):
            self.act[key], self.grad[key] = self._buf(lvl, c), self._buf(lvl, c)

        for k, u in enumerate(U):
            own = self.owns(u.block)
            if own:
                self.raw[u.conv], self.g_raw[u.conv] = self._buf(u.lvl, u.cout), self._buf(u.lvl, u.cout)
                self.bnbuf[u.conv] = torch.zeros(4, u.cout, device=dev)
            consumer_blocks = []
            if k % 2 == 1:  # block output
                if u.block <= 3:
                    consumer_blocks.append(NUM_BLOCKS - 1 - u.block)  # skip tensor
                elif u.block < NUM_BLOCKS - 1:
                    consumer_blocks.append(u.block + 1)
            if own or any(self.owns(cb) for cb in consumer_blocks):
                both(u.conv, u.lvl, u.cout)
        c = self.enc_c
        for l in range(1, 5):  # pooled input of down_l: produced by block l-1 (fused into its last apply pass)
            if self.owns(l - 1) or self.owns(l):
                both(f"pool{l}", l, c[l - 1])
        for i in range(1, 5):
            if self.owns(4 + i):
                both(f"up{i}", 4 - i, self.up_ch(i))
        hb = N.lib().dpt_head_blocks()
        self.head_blocks = hb
        self.head_partial = io["head_partial"] if "head_partial" in io else torch.zeros(hb, 4, device=dev)
        self.loss_state = io["loss_state"] if "loss_state" in io else torch.zeros(8, device=dev)
        self.loss_state[5] = float(self.total_pixels)  # the head's backward reads only this entry
        if self.has_head:
            self.head_scratch = torch.empty(int(N.lib().dpt_head_softmax_scratch_bytes(B * H * W, 64, self.K)), dtype=torch.uint8,
                                            device=dev)

    def _conv_entries(self, oscale: Optional[Dict[str, torch.Tensor]] = None):
        entries = []
        for u in self._owned_units():
            w = self.p(u.conv + ".weight")
            cin_pad = 16 if u.cin < 16 else u.cin
            e = N.PrepEntry(w.data_ptr(), None, self.wf[u.conv].data_ptr(), None if oscale is not None else N.ptr(self.wd[u.conv]),
                            None, u.cout, u.cin, cin_pad, 0)
            if oscale is not None:
                e.oscale = oscale[u.conv][2].data_ptr()
            entries.append(e)
        return entries

    def _build_weights(self) -> None:
        dev = self.device
        self.wf: Dict[str, torch.Tensor] = {}
        self.wd: Dict[str, Optional[torch.Tensor]] = {}
        self.bias4: Dict[str, torch.Tensor] = {}
        for u in self._owned_units():
            cin_pad = 16 if u.cin < 16 else u.cin
            self.wf[u.conv] = torch.zeros(9, u.cout, cin_pad, dtype=torch.bfloat16, device=dev)
            self.wd[u.conv] = torch.zeros(9, u.cin, u.cout, dtype=torch.bfloat16, device=dev) if u.cin >= 16 else None
        entries = self._conv_entries()
        if not self.bilinear:
            for i in range(1, 5):
                if not self.owns(4 + i):
                    continue
                name = f"up{i}.up"
                cin, cout = self.x_ch(i), self.x_ch(i) // 2
                self.wf[name] = torch.zeros(1, 4 * cout, cin, dtype=torch.bfloat16, device=dev)
                self.wd[name] = torch.zeros(4, cin, cout, dtype=torch.bfloat16, device=dev)
                self.bias4[name] = torch.zeros(4 * cout, device=dev)
                entries.append(N.PrepEntry(self.p(name + ".weight").data_ptr(), self.p(name + ".bias").data_ptr(),
                                           self.wf[name].data_ptr(), self.wd[name].data_ptr(), self.bias4[name].data_ptr(), cout,
                                           cin, cin, 1))
        self.prep_table = N.prep_weights(entries, dev) if entries else torch.zeros(1, dtype=torch.uint8, device=dev)
        self.prep_entries = len(entries)
        # evaluation mode: scale / shift from the running statistics; the scale is folded into the forward weights
        self.eval_bn = {u.conv: torch.zeros(4, u.cout, device=dev) for u in self._owned_units()}
        ev = self._conv_entries(self.eval_bn)
        self.eval_table = N.prep_weights(ev, dev) if ev else torch.zeros(1, dtype=torch.uint8, device=dev)

    def refresh_weights(self, stream=None) -> None:
        """Re-derive the bf16 kernel-side weights from the fp32 arena (after load_state_dict / optimizer step / evaluation)."""
        if self.prep_entries:
            N.run_prep_weights(self.prep_table, self.prep_entries, stream)
        self.core_state["weights"] = "train"

    def _build_eval_prep(self, prog: Program) -> None:
        units = self._owned_units()
        for u in units:
            prog.bn_eval_prepare(self.p(u.bn + ".weight"), self.p(u.bn + ".bias"), self._buffer(u.bn + ".running_mean"),
                                 self._buffer(u.bn + ".running_var"), self.eval_bn[u.conv], label=f"eval bn {u.bn}")
        if units:
            prog.prep_weights(self.eval_table, len(units), "prep_weights (BatchNorm folded)")

    def _enter_eval(self, stream=None) -> None:
        if self.core_state["weights"] != "eval":
            core = self.core_state.get("owner", self)
            core.eval_prep.run(0, -1, stream)
            self.core_state["weights"] = "eval"

    def _enter_train(self, stream=None) -> None:
        if self.core_state["weights"] != "train":
            self.refresh_weights(stream)

    # ---- programs ------------------------------------------------------------------------------------------
    def _unit_fwd(self, prog: Program, u: _Unit, srcs, train: bool, pooled: Optional[torch.Tensor] = None) -> torch.Tensor:
        h, w = self.H >> u.lvl, self.W >> u.lvl
        if not train:
            # evaluation: BatchNorm is folded into the weights (scale) and the epilogue (shift as bias, ReLU)
            plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=list(srcs), weight=self.wf[u.conv], cout=u.cout,
                              dsts=[self.act[u.conv]], bias=self.eval_bn[u.conv][3], relu=True)
            prog.conv(plan, f"fwd conv+bn(eval)+relu {u.conv}")
            if pooled is not None:
                prog.maxpool(self.act[u.conv], pooled, f"fwd pool {u.conv}")
            return self.act[u.conv]
        stats = self.bn_scratch.view(torch.float32)
        plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=list(srcs), weight=self.wf[u.conv], cout=u.cout,
                          dsts=[self.raw[u.conv]], stats=stats)
        info = plan.info()
        prog.conv(plan, f"fwd conv+stats {u.conv}")
        gam, bet = self.p(u.bn + ".weight"), self.p(u.bn + ".bias")
        rm, rv = self._buffer(u.bn + ".running_mean"), self._buffer(u.bn + ".running_var")
        n_tiles = u.cout // info["tile_n"]
        prog.bn_finalize(stats, info["grid"] // n_tiles, info["tile_n"], gam, bet, rm, rv, self.bnbuf[u.conv], self.B * h * w,
                         label=f"fwd bn finalize {u.bn}")
        prog.bn_apply(self.raw[u.conv], self.act[u.conv], self.bnbuf[u.conv], pooled,
                      label=f"fwd bn apply+relu{'+pool' if pooled is not None else ''} {u.bn}")
        return self.act[u.conv]

    def _build_forward(self, prog: Program, train: bool) -> None:
        U = self._units()
        a = self.act
        # op index range of every owned block (pipeline stages run, wait and send block by block)
        ranges: Dict[int, Tuple[int, int]] = {}
        if train:
            self.fwd_block_range = ranges
        else:
            self.infer_block_range = ranges
        mark = len(prog)

        def close(block: int) -> None:
            nonlocal mark
            ranges[block] = (mark, len(prog))
            mark = len(prog)

        if self.has_input:
            prog.pack_input(self.images, self.x0)
            if train:
                prog.record(EV_IMAGES_FREE, label="input images consumed")
        for l in range(0, 5):
            if not self.owns(l):
                continue
            x = self.x0 if l == 0 else a[f"pool{l}"]
            x = self._unit_fwd(prog, U[2 * l], [x], train)
            self._unit_fwd(prog, U[2 * l + 1], [x], train, pooled=a[f"pool{l + 1}"] if l < 4 else None)
            close(l)
        for i in range(1, 5):
            if not self.owns(4 + i):
                continue
            lvl = 4 - i
            x = a[U[9].conv] if i == 1 else a[U[11 + 2 * (i - 2)].conv]
            if self.bilinear:
                prog.bilinear_fwd(x, a[f"up{i}"], f"fwd bilinear up{i}")
            else:
                name = f"up{i}.up"
                plan = N.ConvPlan(mode=N.CONV_1X1, out_mode=N.OUT_SHUFFLE2, batch=self.B, h=self.H >> (lvl + 1),
                                  w=self.W >> (lvl + 1), srcs=[x], weight=self.wf[name], cout=4 * a[f"up{i}"].shape[-1],
                                  dsts=[a[f"up{i}"]], bias=self.bias4[name])
                prog.conv(plan, f"fwd convT up{i}")
            x = self._unit_fwd(prog, U[10 + 2 * (i - 1)], [a[U[2 * lvl + 1].conv], a[f"up{i}"]], train)
            self._unit_fwd(prog, U[11 + 2 * (i - 1)], [x], train)
            if i < 4:
                close(4 + i)
        if not self.has_head:
            return
        x = a[U[-1].conv]
        w, b = self.p("outc.conv.weight"), self.p("outc.conv.bias")
        if not train:
            prog.head_softmax_fwd(x, w, b, None, self.target_dtype, self.logits, self.head_partial, None, self.K, "fwd outc")
            close(NUM_BLOCKS - 1)
            return
        single = self.pre_finalize is None and self.finalize
        prog.head_softmax_fwd(x, w, b, self.target, self.target_dtype, self.logits, self.head_partial,
                              self.loss_state if single else None, self.K, "fwd outc+softmax+loss")
        if self.pre_finalize is not None:  # e.g. DataParallel: all-reduce the loss statistics across replicas
            self.pre_finalize(prog)
        if self.finalize and not single:
            prog.head_softmax_finalize([self.head_partial], self.total_pixels, self.loss_state, "loss finalize")
        close(NUM_BLOCKS - 1)

    def _unit_bwd(self, prog: Program, u: _Unit, srcs, dsts) -> None:
        """grad[u] holds d(loss)/d(act); produces parameter gradients and the gradients of the sources in `dsts`."""
        h, w = self.H >> u.lvl, self.W >> u.lvl
        prog.bn_bwd(self.raw[u.conv], self.grad[u.conv], self.g_raw[u.conv], self.bnbuf[u.conv], self.g(u.bn + ".weight"),
                    self.g(u.bn + ".bias"), self.bn_scratch, accumulate=self.accumulate, label=f"bwd bn {u.bn}")
        cin = sum(s.shape[-1] for s in srcs)
        n_valid = 3 if u.cin == 3 else 0
        wp = N.WgradPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, z=self.g_raw[u.conv], xs=list(srcs),
                         dw=self.g(u.conv + ".weight"), strides=(u.cin * 9, 9, 1), scratch=self.wgrad_scratch, n_valid=n_valid,
                         accumulate=self.accumulate)
        plan = None
        if dsts is not None:
            plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=[self.g_raw[u.conv]], weight=self.wd[u.conv], cout=cin,
                              dsts=list(dsts))
        self._wgrad_dgrad(prog, wp, f"bwd wgrad {u.conv}", u.conv, plan, f"bwd dgrad {u.conv}")

    def _wgrad_dgrad(self, prog: Program, wp, wlabel: str, name: str, dgrad, dlabel: str) -> None:
        """The two consumers of an output gradient the main stream has just produced.  Two-stream backward: the data
        gradient (critical path) is enqueued first on the main stream, the weight gradient on the second stream."""
        if self.wgrad_concurrent:
            prog.record(EV_DZ_READY, label="output gradient ready")
            if dgrad is not None:
                prog.conv(dgrad, dlabel)
            prog.wait(EV_DZ_READY, stream=1, label="wgrad stream: output gradient ready")
            prog.wgrad(wp, wlabel, stream=1)
            return
        prog.wgrad(wp, wlabel)
        if self.grads_ready_hook is not None:
            self.grads_ready_hook(self, prog, name)
        if dgrad is not None:
            prog.conv(dgrad, dlabel)

    def _wgrad_scratch_needed(self) -> int:
        need = 16
        for u in self._owned_units():
            h, w = self.H >> u.lvl, self.W >> u.lvl
            if u.conv.startswith("up") and u.conv.endswith(".0"):
                i = 4 - u.lvl
                xcs = [self.enc_c[u.lvl], self.up_ch(i)]
            else:
                xcs = [max(u.cin, 16)]
            need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_3X3, batch=self.B, h=h, w=w, z_c=u.cout, x_cs=xcs))
        if not self.bilinear:
            for i in range(1, 5):
                if not self.owns(4 + i):
                    continue
                lvl = 4 - i
                need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_2X2S2, batch=self.B, h=self.H >> (lvl + 1),
                                                               w=self.W >> (lvl + 1), z_c=self.x_ch(i), x_cs=[self.x_ch(i) // 2]))
        return need

    def _build_backward(self, prog: Program) -> None:
        U = self._units()
        a, g = self.act, self.grad
        ranges = self.bwd_block_range = {}
        mark = len(prog)

        def close(block: int) -> None:
            nonlocal mark
            ranges[block] = (mark, len(prog))
            mark = len(prog)

        if self.has_head:
            last = U[-1]
            prog.head_softmax_bwd(a[last.conv], self.p("outc.conv.weight"), self.p("outc.conv.bias"), self.target,
                                  self.target_dtype, self.loss_state, g[last.conv], self.g("outc.conv.weight"),
                                  self.g("outc.conv.bias"), self.head_scratch, self.K, self.loss_scale, accumulate=self.accumulate,
                                  label="bwd outc+softmax")
            prog.record(EV_TARGET_FREE, label="target consumed")
            if self.grads_ready_hook is not None:
                self.grads_ready_hook(self, prog, "outc")
        for i in range(4, 0, -1):
            if not self.owns(4 + i):
                continue
            lvl = 4 - i
            u0, u1 = U[10 + 2 * (i - 1)], U[11 + 2 * (i - 1)]
            skip = U[2 * lvl + 1].conv
            self._unit_bwd(prog, u1, [a[u0.conv]], [g[u0.conv]])
            self._unit_bwd(prog, u0, [a[skip], a[f"up{i}"]], [g[skip], g[f"up{i}"]])
            prev = U[9].conv if i == 1 else U[11 + 2 * (i - 2)].conv
            if self.bilinear:
                prog.bilinear_bwd(g[f"up{i}"], g[prev], f"bwd bilinear up{i}")
            else:
                name = f"up{i}.up"
                hh, ww = self.H >> (lvl + 1), self.W >> (lvl + 1)
                cin, cout = a[prev].shape[-1], a[f"up{i}"].shape[-1]
                wp = N.WgradPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, z=a[prev], xs=[g[f"up{i}"]],
                                 dw=self.g(name + ".weight"), strides=(cout * 4, 4, 1), scratch=self.wgrad_scratch,
                                 accumulate=self.accumulate, dbias=self.g(name + ".bias"))
                plan = N.ConvPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, srcs=[g[f"up{i}"]], weight=self.wd[name], cout=cin,
                                  dsts=[g[prev]])
                self._wgrad_dgrad(prog, wp, f"bwd wgrad+bgrad {name}", name, plan, f"bwd dgrad {name}")
            close(4 + i)
        # encoder: level 4 has no skip consumer (its activation only feeds up1); levels 3..0 add the pool gradient
        for l in range(4, -1, -1):
            if not self.owns(l):
                continue
            u0, u1 = U[2 * l], U[2 * l + 1]
            if l < 4:
                # grad[u1] holds the skip gradient; add the routed pool gradient (the ReLU mask is applied by bn_bwd)
                prog.pool_bwd(a[u1.conv], g[f"pool{l + 1}"], g[u1.conv], g[u1.conv], False, f"bwd pool{l + 1}+skip")
            self._unit_bwd(prog, u1, [a[u0.conv]], [g[u0.conv]])
            if l > 0:
                self._unit_bwd(prog, u0, [a[f"pool{l}"]], [g[f"pool{l}"]])
            else:
                self._unit_bwd(prog, u0, [self.x0], None)
            if l == 0 and self.wgrad_concurrent:  # join: every weight gradient is complete before the update reads them
                prog.record(EV_WGRADS_DONE, stream=1, label="weight gradients done")
                prog.wait(EV_WGRADS_DONE, label="join the wgrad stream")
            close(l)

    def _build_update(self, prog: Program) -> None:
        if self.numel == 0:
            return
        prog.opt_tick(self.opt_state)
        if self.optimizer == "rmsprop":
            prog.rmsprop(self.params, self.grads, self.s0, self.s1 if self.momentum > 0 else None, self.numel, self.opt_state,
                         self.alpha, self.eps, self.weight_decay, self.momentum)
        else:
            prog.adam(self.params, self.grads, self.s0, self.s1, self.numel, self.opt_state, self.eps, self.weight_decay)
        prog.prep_weights(self.prep_table, self.prep_entries)

    # ---- API -----------------------------------------------------------------------------------------------
    # optimizer-state names shared with NativeUNetEngine (DataParallel / DDP descriptors)
    @property
    def m(self):
        return self.s0

    @property
    def v(self):
        return self.s1

    @property
    def rms_alpha(self):
        return self.alpha

    @property
    def rms_momentum(self):
        return self.momentum

    def set_lr(self, lr: float) -> None:
        self.lr = lr
        with torch.cuda.device(self.device):
            N.check(N.lib().dpt_opt_set_lr(self.opt_state.data_ptr(), lr, N.stream_ptr()))

    def load_batch(self, images, masks) -> None:
        if images is not None:
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
        if masks is not None:
            self.target.copy_(masks.reshape(self.target.shape), non_blocking=True)
        self._input_dirty = True

    @property
    def launches_per_step(self) -> int:
        extra = {"bwd wgrad": 1, "bwd bn": 2, "bwd outc": 2, "fwd outc+softmax": 1}
        from .program import OP_EVENT_RECORD, OP_STREAM_WAIT

        n = 0
        for lbl, kind in zip(self.fwd.labels[:self.step_end], self.fwd.kinds[:self.step_end]):
            if kind not in (OP_EVENT_RECORD, OP_STREAM_WAIT):
                n += 1 + next((v for k, v in extra.items() if lbl.startswith(k)), 0)
        return n

    def prefetch(self, images, masks) -> None:
        """Copy the NEXT batch straight into the input buffers on a side stream while the current step computes; the copy
        waits for the program events "images consumed" / "target consumed" of the step in flight (see NativeUNetEngine)."""
        if not hasattr(self, "_copy_stream"):
            with torch.cuda.device(self.device):
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._staged = torch.cuda.Event()
        cs = self._copy_stream
        if getattr(self, "_input_dirty", False):
            cs.wait_stream(torch.cuda.current_stream(self.device))
            self._input_dirty = False
        with torch.cuda.stream(cs):
            self.fwd.stream_wait(EV_IMAGES_FREE, cs)
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
            self.fwd.stream_wait(EV_TARGET_FREE, cs)
            self.target.copy_(masks.reshape(self.target.shape), non_blocking=True)
            self._staged.record(cs)
        self._has_staged = True

    def _consume_staged(self) -> None:
        torch.cuda.current_stream(self.device).wait_event(self._staged)
        self._has_staged = False

    def _train_step_partial(self, images, masks, stream) -> torch.Tensor:
        b = int(images.shape[0])
        if not self.supports_partial_batch or not 0 < b < self.B:
            raise ValueError(f"batch of {b} images on an engine built for {self.B}")
        eng = self._partial.get(b)
        with torch.cuda.device(self.device):
            if eng is None:
                eng = NativeMilesialEngine(self.model, b, self.H, self.W, self.device, shared=self, build_update=False,
                                           **self._partial_kwargs)
                self._partial[b] = eng
            eng.load_batch(images, masks)
            self._enter_train(stream)
            eng.fwd.run(0, eng.bwd_end, stream)
            self.fwd.run(self.bwd_end, self.step_end, stream, getattr(self, "aux_stream", None))
        return eng.loss_state[4]

    def train_step(self, images=None, masks=None, stream=None) -> torch.Tensor:
        if images is not None and images.shape[0] != self.B:
            return self._train_step_partial(images, masks, stream)
        with torch.cuda.device(self.device):
            if images is not None:
                self.load_batch(images, masks)
            elif getattr(self, "_has_staged", False):
                self._consume_staged()
            self._enter_train(stream)
            self.fwd.run(0, self.step_end, stream, getattr(self, "aux_stream", None))
            self._input_dirty = False
        return self.loss_state[4]

    def forward_backward(self, images=None, masks=None, stream=None) -> torch.Tensor:
        with torch.cuda.device(self.device):
            if images is not None:
                self.load_batch(images, masks)
            self._enter_train(stream)
            self.fwd.run(0, self.bwd_end, stream)
        return self.loss_state[4]

    def run_forward(self, stream=None) -> None:
        self._enter_train(stream)
        self.fwd.run(0, self.fwd_end, stream)

    def run_backward(self, stream=None) -> None:
        self.fwd.run(self.fwd_end, self.bwd_end, stream)

    def run_update(self, stream=None) -> None:
        self.fwd.run(self.bwd_end, self.step_end, stream)

    def run_eval_forward(self, stream=None) -> None:
        self._enter_eval(stream)
        self.evalp.run(0, -1, stream)

    def evaluate_batch(self, images, masks, stream=None) -> torch.Tensor:
        """Loss in evaluation mode (running statistics), like evaluate.py does after model.eval()."""
        with torch.cuda.device(self.device):
            self.load_batch(images, masks)
            self.run_eval_forward(stream)
            # the eval program has no target: compute the loss from the logits with the torch criterion (rare path)
            from ..utils.loss import CEDiceLoss

            return CEDiceLoss(self.K)(self.logits, self.target.long())

    def forward(self, images, stream=None) -> torch.Tensor:
        """Logits [B,K,H,W] fp32 in evaluation mode."""
        with torch.cuda.device(self.device):
            self.load_batch(images, None)
            self.run_eval_forward(stream)
        return self.logits

    def optimizer_state(self) -> dict:
        return {"m": self.s0.detach().cpu(), "v": self.s1.detach().cpu(), "opt_state": self.opt_state.detach().cpu()}

    def load_optimizer_state(self, sd: dict) -> None:
        self.s0.copy_(sd["m"])
        self.s1.copy_(sd["v"])
        self.opt_state.copy_(sd["opt_state"])
