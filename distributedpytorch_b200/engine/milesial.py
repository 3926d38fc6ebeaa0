"""Native engine for the north-star model family: `UNet(variant='milesial', n_classes>=2, bilinear=False|True)`
(DoubleConv / Down / Up / OutConv with BatchNorm, model/modelsummary.txt:153-247 of the reference).

Same construction as engine/native_unet.py - one flat fp32 parameter arena with the module's parameters as views, a
native launch program for forward / loss / hand-written backward / optimizer - with the extra pieces of this family:

  conv (tcgen05, no bias) -> BatchNorm batch statistics (deterministic two-stage) -> normalise + ReLU
  max-pool -> DoubleConv;  ConvTranspose (GEMM + pixel shuffle) or bilinear x2;  concat fused into the conv K loop
  OutConv + softmax + cross-entropy + Dice in one pass;  RMSprop (weight_decay 1e-8, momentum 0.999) or Adam
  backward: BatchNorm backward recomputes the ReLU mask from the saved pre-norm activation, so the data-gradient
  convolutions run without a mask; the pool backward adds the skip gradient.
BatchNorm running statistics live in the module's buffers (device tensors) and are updated in place by the
forward kernel; `eval_forward` uses them.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Tuple

import torch

from ..models.unet import UNet
from ..ops import native as N
from .program import Program


class _Unit:
    """One Conv3x3(bias=False) -> BatchNorm -> ReLU unit."""

    def __init__(self, conv_name: str, bn_name: str, cin: int, cout: int, lvl: int):
        self.conv, self.bn, self.cin, self.cout, self.lvl = conv_name, bn_name, cin, cout, lvl


class NativeMilesialEngine:
    def __init__(self, model: UNet, batch: int, height: int, width: int, device, *, lr: float = 1e-5,
                 weight_decay: float = 1e-8, optimizer: str = "rmsprop", momentum: float = 0.999, alpha: float = 0.99,
                 eps: float = 1e-8, betas=(0.9, 0.999), loss_scale: float = 1.0, params: Optional[torch.Tensor] = None,
                 grads: Optional[torch.Tensor] = None, build_update: bool = True):
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
        f = 2 if self.bilinear else 1
        self.enc_c = [64, 128, 256, 512, 1024 // f]          # inc, down1..down4 output channels
        self.dec_out = [512 // f, 256 // f, 128 // f, 64]     # up1..up4 output channels
        N.lib()
        with torch.cuda.device(self.device):
            for b in model.buffers():
                b.data = b.data.to(self.device)
            self._layout_params(params, grads)
            self._alloc()
            self._build_weights()
            self.fwd = Program(self.device)
            self._build_forward(self.fwd, train=True)
            self.fwd_end = len(self.fwd)
            self._build_backward(self.fwd)
            self.bwd_end = len(self.fwd)
            if build_update:
                self._build_update(self.fwd)
            self.step_end = len(self.fwd)
            self.evalp = Program(self.device)
            self._build_forward(self.evalp, train=False)
            self.refresh_weights()

    # ---- structure ----------------------------------------------------------------------------------------
    def _units(self) -> List[_Unit]:
        U: List[_Unit] = []
        c = self.enc_c
        U += [_Unit("inc.double_conv.0", "inc.double_conv.1", 3, 64, 0), _Unit("inc.double_conv.3", "inc.double_conv.4", 64, 64, 0)]
        for l in range(1, 5):
            base = f"down{l}.maxpool_conv.1.double_conv"
            U += [_Unit(base + ".0", base + ".1", c[l - 1], c[l], l), _Unit(base + ".3", base + ".4", c[l], c[l], l)]
        for i in range(1, 5):
            lvl = 4 - i
            cin = c[lvl] + self.up_ch(i)
            cout = self.dec_out[i - 1]
            mid = cout if not self.bilinear else cin // 2
            base = f"up{i}.conv.double_conv"
            U += [_Unit(base + ".0", base + ".1", cin, mid, lvl), _Unit(base + ".3", base + ".4", mid, cout, lvl)]
        return U

    def x_ch(self, i: int) -> int:
        """Channels of the tensor entering Up block i (down4's output, then the previous Up block's output)."""
        return self.enc_c[4] if i == 1 else self.dec_out[i - 2]

    def up_ch(self, i: int) -> int:
        """Channels after the up-sampling of Up block i: ConvTranspose halves them, bilinear keeps them."""
        return self.x_ch(i) if self.bilinear else self.x_ch(i) // 2

    def _layout_params(self, params, grads) -> None:
        have = dict(self.model.named_parameters())
        order: List[str] = []
        for u in self._units()[:10]:
            order += [u.conv + ".weight", u.bn + ".weight", u.bn + ".bias"]
        units = self._units()
        for i in range(1, 5):
            if not self.bilinear:
                order += [f"up{i}.up.weight", f"up{i}.up.bias"]
            for u in units[10 + 2 * (i - 1):12 + 2 * (i - 1)]:
                order += [u.conv + ".weight", u.bn + ".weight", u.bn + ".bias"]
        order += ["outc.conv.weight", "outc.conv.bias"]
        assert set(order) == set(have.keys()), sorted(set(have.keys()) ^ set(order))[:5]
        self.slots: Dict[str, Tuple[int, int, Tuple[int, ...]]] = {}
        off = 0
        for n in order:
            p = have[n]
            self.slots[n] = (off, p.numel(), tuple(p.shape))
            off += (p.numel() + 3) // 4 * 4
        self.numel = off
        dev = self.device
        self.params = params if params is not None else torch.zeros(off, device=dev)
        self.grads = grads if grads is not None else torch.zeros(off, device=dev)
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

    def _alloc(self) -> None:
        dev, B, H, W = self.device, self.B, self.H, self.W
        self.images = torch.zeros(B, 3, H, W, device=dev)
        self.target = torch.zeros(B, H, W, dtype=torch.uint8, device=dev)
        self.logits = torch.zeros(B, self.K, H, W, device=dev)
        self.x0 = self._buf(0, 16)
        self.raw: Dict[str, torch.Tensor] = {}   # pre-norm conv outputs
        self.act: Dict[str, torch.Tensor] = {}   # post BN+ReLU
        self.g_act: Dict[str, torch.Tensor] = {}  # gradient w.r.t. act
        self.g_raw: Dict[str, torch.Tensor] = {}  # gradient w.r.t. raw (dz of the conv)
        self.bnbuf: Dict[str, torch.Tensor] = {}
        for u in self._units():
            self.raw[u.conv], self.act[u.conv] = self._buf(u.lvl, u.cout), self._buf(u.lvl, u.cout)
            self.g_act[u.conv], self.g_raw[u.conv] = self._buf(u.lvl, u.cout), self._buf(u.lvl, u.cout)
            self.bnbuf[u.conv] = torch.zeros(4, u.cout, device=dev)
        c = self.enc_c
        self.pool = {l: self._buf(l, c[l - 1]) for l in range(1, 5)}      # pooled input of down_l
        self.g_pool = {l: self._buf(l, c[l - 1]) for l in range(1, 5)}
        self.up: Dict[int, torch.Tensor] = {}
        self.g_up: Dict[int, torch.Tensor] = {}
        for i in range(1, 5):
            lvl = 4 - i
            self.up[i], self.g_up[i] = self._buf(lvl, self.up_ch(i)), self._buf(lvl, self.up_ch(i))
        hb = N.lib().dpt_head_blocks()
        self.head_partial = torch.zeros(hb, 4, device=dev)
        self.loss_state = torch.zeros(8, device=dev)
        self.opt_state = torch.zeros(8, device=dev)
        self.s0 = torch.zeros(self.numel, device=dev)
        self.s1 = torch.zeros(self.numel, device=dev)
        L = N.lib()
        self.bn_scratch = torch.empty(int(L.dpt_bn_scratch_bytes(1024)) + 2 * 1024 * 4, dtype=torch.uint8, device=dev)
        self.head_scratch = torch.empty(int(L.dpt_head_softmax_scratch_bytes(B * H * W, 64, self.K)), dtype=torch.uint8, device=dev)
        self.csum_scratch = torch.empty(int(L.dpt_channel_sum_scratch_bytes(1024)), dtype=torch.uint8, device=dev)
        N.check(L.dpt_opt_init(self.opt_state.data_ptr(), self.lr, 0, self.betas[0], self.betas[1], N.stream_ptr()))

    def _build_weights(self) -> None:
        dev = self.device
        self.wf: Dict[str, torch.Tensor] = {}
        self.wd: Dict[str, Optional[torch.Tensor]] = {}
        self.bias4: Dict[str, torch.Tensor] = {}
        entries = []
        for u in self._units():
            w = self.p(u.conv + ".weight")
            cin_pad = 16 if u.cin < 16 else u.cin
            self.wf[u.conv] = torch.zeros(9, u.cout, cin_pad, dtype=torch.bfloat16, device=dev)
            self.wd[u.conv] = torch.zeros(9, u.cin, u.cout, dtype=torch.bfloat16, device=dev) if u.cin >= 16 else None
            entries.append(N.PrepEntry(w.data_ptr(), None, self.wf[u.conv].data_ptr(), N.ptr(self.wd[u.conv]), None, u.cout, u.cin,
                                       cin_pad, 0))
        if not self.bilinear:
            for i in range(1, 5):
                name = f"up{i}.up"
                cin, cout = self.x_ch(i), self.x_ch(i) // 2
                self.wf[name] = torch.zeros(1, 4 * cout, cin, dtype=torch.bfloat16, device=dev)
                self.wd[name] = torch.zeros(4, cin, cout, dtype=torch.bfloat16, device=dev)
                self.bias4[name] = torch.zeros(4 * cout, device=dev)
                entries.append(N.PrepEntry(self.p(name + ".weight").data_ptr(), self.p(name + ".bias").data_ptr(),
                                           self.wf[name].data_ptr(), self.wd[name].data_ptr(), self.bias4[name].data_ptr(), cout,
                                           cin, cin, 1))
        self.prep_table = N.prep_weights(entries, dev)
        self.prep_entries = len(entries)

    def refresh_weights(self, stream=None) -> None:
        N.run_prep_weights(self.prep_table, self.prep_entries, stream)

    # ---- programs ------------------------------------------------------------------------------------------
    def _unit_fwd(self, prog: Program, u: _Unit, srcs, train: bool) -> torch.Tensor:
        h, w = self.H >> u.lvl, self.W >> u.lvl
        plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=list(srcs), weight=self.wf[u.conv], cout=u.cout,
                          dsts=[self.raw[u.conv]])
        prog.conv(plan, f"fwd conv {u.conv}")
        gam, bet = self.p(u.bn + ".weight"), self.p(u.bn + ".bias")
        rm, rv = self._buffer(u.bn + ".running_mean"), self._buffer(u.bn + ".running_var")
        if train:
            prog.bn_fwd_train(self.raw[u.conv], self.act[u.conv], gam, bet, rm, rv, self.bnbuf[u.conv], self.bn_scratch,
                              label=f"fwd bn+relu {u.bn}")
        else:
            prog.bn_fwd_eval(self.raw[u.conv], self.act[u.conv], gam, bet, rm, rv, self.bnbuf[u.conv],
                             label=f"fwd bn(eval)+relu {u.bn}")
        return self.act[u.conv]

    def _build_forward(self, prog: Program, train: bool) -> None:
        U = self._units()
        prog.pack_input(self.images, self.x0)
        x = self._unit_fwd(prog, U[0], [self.x0], train)
        x = self._unit_fwd(prog, U[1], [x], train)
        self.skips = [x]
        for l in range(1, 5):
            prog.maxpool(x, self.pool[l], f"fwd pool{l}")
            x = self._unit_fwd(prog, U[2 * l], [self.pool[l]], train)
            x = self._unit_fwd(prog, U[2 * l + 1], [x], train)
            self.skips.append(x)
        for i in range(1, 5):
            lvl = 4 - i
            if self.bilinear:
                prog.bilinear_fwd(x, self.up[i], f"fwd bilinear up{i}")
            else:
                name = f"up{i}.up"
                plan = N.ConvPlan(mode=N.CONV_1X1, out_mode=N.OUT_SHUFFLE2, batch=self.B, h=self.H >> (lvl + 1),
                                  w=self.W >> (lvl + 1), srcs=[x], weight=self.wf[name], cout=4 * self.up[i].shape[-1],
                                  dsts=[self.up[i]], bias=self.bias4[name])
                prog.conv(plan, f"fwd convT up{i}")
            x = self._unit_fwd(prog, U[10 + 2 * (i - 1)], [self.skips[lvl], self.up[i]], train)
            x = self._unit_fwd(prog, U[11 + 2 * (i - 1)], [x], train)
        self.final_act = x
        w, b = self.p("outc.conv.weight"), self.p("outc.conv.bias")
        prog.head_softmax_fwd(x, w, b, self.target if train else None, self.target_dtype, self.logits, self.head_partial,
                              self.loss_state, self.K, "fwd outc+softmax+loss" if train else "fwd outc")

    def _unit_bwd(self, prog: Program, u: _Unit, srcs, dsts) -> None:
        """g_act[u] holds d(loss)/d(act); produces parameter gradients and the gradients of the sources in `dsts`."""
        h, w = self.H >> u.lvl, self.W >> u.lvl
        prog.bn_bwd(self.raw[u.conv], self.g_act[u.conv], self.g_raw[u.conv], self.bnbuf[u.conv], self.g(u.bn + ".weight"),
                    self.g(u.bn + ".bias"), self.bn_scratch, label=f"bwd bn {u.bn}")
        cin = sum(s.shape[-1] for s in srcs)
        n_valid = 3 if u.cin == 3 else 0
        cin_real = u.cin
        wp = N.WgradPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, z=self.g_raw[u.conv], xs=list(srcs),
                         dw=self.g(u.conv + ".weight"), strides=(cin_real * 9, 9, 1), scratch=self.wgrad_scratch, n_valid=n_valid)
        prog.wgrad(wp, f"bwd wgrad {u.conv}")
        if dsts is not None:
            plan = N.ConvPlan(mode=N.CONV_3X3, batch=self.B, h=h, w=w, srcs=[self.g_raw[u.conv]], weight=self.wd[u.conv], cout=cin,
                              dsts=list(dsts))
            prog.conv(plan, f"bwd dgrad {u.conv}")

    def _wgrad_scratch_needed(self) -> int:
        need = 16
        for u in self._units():
            h, w = self.H >> u.lvl, self.W >> u.lvl
            if u.conv.startswith("up") and u.conv.endswith(".0"):
                lvl = u.lvl
                i = 4 - lvl
                xcs = [self.skips_c[lvl], self.up_c[i]]
            else:
                xcs = [max(u.cin, 16)]
            need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_3X3, batch=self.B, h=h, w=w, z_c=u.cout, x_cs=xcs))
        if not self.bilinear:
            for i in range(1, 5):
                lvl = 4 - i
                need = max(need, N.WgradPlan.scratch_bytes_for(mode=N.CONV_2X2S2, batch=self.B, h=self.H >> (lvl + 1),
                                                               w=self.W >> (lvl + 1), z_c=self.x_ch(i), x_cs=[self.x_ch(i) // 2]))
        return need

    def _build_backward(self, prog: Program) -> None:
        U = self._units()
        self.skips_c = {l: self.enc_c[l] for l in range(5)}
        self.up_c = {i: self.up[i].shape[-1] for i in range(1, 5)}
        self.wgrad_scratch = torch.empty(self._wgrad_scratch_needed(), dtype=torch.uint8, device=self.device)
        last = U[-1]
        prog.head_softmax_bwd(self.final_act, self.p("outc.conv.weight"), self.p("outc.conv.bias"), self.target, self.target_dtype,
                              self.loss_state, self.g_act[last.conv], self.g("outc.conv.weight"), self.g("outc.conv.bias"),
                              self.head_scratch, self.K, self.loss_scale, label="bwd outc+softmax")
        for i in range(4, 0, -1):
            lvl = 4 - i
            u0, u1 = U[10 + 2 * (i - 1)], U[11 + 2 * (i - 1)]
            self._unit_bwd(prog, u1, [self.act[u0.conv]], [self.g_act[u0.conv]])
            skip_unit = U[2 * lvl + 1]
            self._unit_bwd(prog, u0, [self.skips[lvl], self.up[i]], [self.g_skip(lvl), self.g_up[i]])
            prev_unit = U[9] if i == 1 else U[11 + 2 * (i - 2)]
            prev_act, g_prev = self.act[prev_unit.conv], self.g_act[prev_unit.conv]
            if self.bilinear:
                prog.bilinear_bwd(self.g_up[i], g_prev, f"bwd bilinear up{i}")
            else:
                name = f"up{i}.up"
                hh, ww = self.H >> (lvl + 1), self.W >> (lvl + 1)
                cin, cout = prev_act.shape[-1], self.up[i].shape[-1]
                wp = N.WgradPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, z=prev_act, xs=[self.g_up[i]],
                                 dw=self.g(name + ".weight"), strides=(cout * 4, 4, 1), scratch=self.wgrad_scratch)
                prog.wgrad(wp, f"bwd wgrad {name}")
                prog.channel_sum(self.g_up[i], self.g(name + ".bias"), self.csum_scratch, 1.0, f"bwd bgrad {name}")
                plan = N.ConvPlan(mode=N.CONV_2X2S2, batch=self.B, h=hh, w=ww, srcs=[self.g_up[i]], weight=self.wd[name], cout=cin,
                                  dsts=[g_prev])
                prog.conv(plan, f"bwd dgrad {name}")
            del skip_unit
        # encoder: level 4 has no skip consumer (its activation only feeds up1); levels 3..0 add the pool gradient
        for l in range(4, 0, -1):
            u0, u1 = U[2 * l], U[2 * l + 1]
            if l < 4:
                # g_act[u1] currently holds the skip gradient; add the routed pool gradient (mask is applied by bn_bwd)
                prog.pool_bwd(self.act[u1.conv], self.g_pool[l + 1], self.g_act[u1.conv], self.g_act[u1.conv], False,
                              f"bwd pool{l + 1}+skip")
            self._unit_bwd(prog, u1, [self.act[u0.conv]], [self.g_act[u0.conv]])
            self._unit_bwd(prog, u0, [self.pool[l]], [self.g_pool[l]])
        u0, u1 = U[0], U[1]
        prog.pool_bwd(self.act[u1.conv], self.g_pool[1], self.g_act[u1.conv], self.g_act[u1.conv], False, "bwd pool1+skip")
        self._unit_bwd(prog, u1, [self.act[u0.conv]], [self.g_act[u0.conv]])
        self._unit_bwd(prog, u0, [self.x0], None)

    def g_skip(self, lvl: int) -> torch.Tensor:
        """Gradient buffer of the skip activation of level `lvl` (= g_act of the encoder unit that produced it)."""
        return self.g_act[self._units()[2 * lvl + 1].conv]

    def _build_update(self, prog: Program) -> None:
        prog.opt_tick(self.opt_state)
        if self.optimizer == "rmsprop":
            prog.rmsprop(self.params, self.grads, self.s0, self.s1 if self.momentum > 0 else None, self.numel, self.opt_state,
                         self.alpha, self.eps, self.weight_decay, self.momentum)
        else:
            prog.adam(self.params, self.grads, self.s0, self.s1, self.numel, self.opt_state, self.eps, self.weight_decay)
        prog.prep_weights(self.prep_table, self.prep_entries)

    # ---- API -----------------------------------------------------------------------------------------------
    def set_lr(self, lr: float) -> None:
        self.lr = lr
        N.check(N.lib().dpt_opt_set_lr(self.opt_state.data_ptr(), lr, N.stream_ptr()))

    def load_batch(self, images, masks) -> None:
        if images is not None:
            self.images.copy_(images.reshape(self.images.shape), non_blocking=True)
        if masks is not None:
            self.target.copy_(masks.reshape(self.target.shape), non_blocking=True)

    @property
    def launches_per_step(self) -> int:
        extra = {"bwd wgrad": 1, "bwd bgrad": 1, "fwd bn": 2, "bwd bn": 2, "bwd outc": 2, "fwd outc": 1}
        n = 0
        for lbl in self.fwd.labels[:self.step_end]:
            n += 1 + next((v for k, v in extra.items() if lbl.startswith(k)), 0)
        return n

    def prefetch(self, images, masks) -> None:
        """Stage the next batch on a side stream while the current step computes (see NativeUNetEngine.prefetch)."""
        if not hasattr(self, "_copy_stream"):
            with torch.cuda.device(self.device):
                self._copy_stream = torch.cuda.Stream(device=self.device)
                self._stage_images, self._stage_target = torch.empty_like(self.images), torch.empty_like(self.target)
                self._staged, self._consumed = torch.cuda.Event(), torch.cuda.Event()
                self._consumed.record(torch.cuda.current_stream(self.device))
        self._copy_stream.wait_event(self._consumed)
        with torch.cuda.stream(self._copy_stream):
            self._stage_images.copy_(images.reshape(self.images.shape), non_blocking=True)
            self._stage_target.copy_(masks.reshape(self.target.shape), non_blocking=True)
            self._staged.record(self._copy_stream)
        self._has_staged = True

    def train_step(self, images=None, masks=None, stream=None) -> torch.Tensor:
        if images is not None:
            self.load_batch(images, masks)
        elif getattr(self, "_has_staged", False):
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(self._staged)
            self.images.copy_(self._stage_images, non_blocking=True)
            self.target.copy_(self._stage_target, non_blocking=True)
            self._consumed.record(cur)
            self._has_staged = False
        self.fwd.run(0, self.step_end, stream)
        return self.loss_state[4]

    def forward_backward(self, images=None, masks=None, stream=None) -> torch.Tensor:
        if images is not None:
            self.load_batch(images, masks)
        self.fwd.run(0, self.bwd_end, stream)
        return self.loss_state[4]

    def evaluate_batch(self, images, masks, stream=None) -> torch.Tensor:
        """Loss in evaluation mode (running statistics), like evaluate.py does after model.eval()."""
        self.load_batch(images, masks)
        self.evalp.run(0, -1, stream)
        # the eval program has no target: compute the loss from the logits with the torch criterion (rare path)
        from ..utils.loss import CEDiceLoss

        return CEDiceLoss(self.K)(self.logits, self.target.long())

    def forward(self, images, stream=None) -> torch.Tensor:
        """Logits [B,K,H,W] fp32 in evaluation mode."""
        self.load_batch(images, None)
        self.evalp.run(0, -1, stream)
        return self.logits
