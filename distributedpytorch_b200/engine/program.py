"""Python builder for native launch programs (csrc/runtime/program.cu).

`Program` collects op records; every helper mirrors one C launcher of include/dpt_b200.h.  The tensors and
plans an op points at are kept alive by the program."""
from __future__ import annotations

import ctypes as C
import os
from typing import List, Optional, Sequence

import torch

from ..ops import native as N

(OP_CONV, OP_WGRAD, OP_PACK_INPUT, OP_MAXPOOL, OP_POOL_BWD, OP_CHANNEL_SUM, OP_PREP_WEIGHTS, OP_HEAD_FWD,
 OP_HEAD_FINALIZE, OP_HEAD_BWD, OP_OPT_TICK, OP_ADAM, OP_RMSPROP, OP_FUSED_AR, OP_BCAST, OP_ALLREDUCE_SUM, OP_MEMSET,
 OP_MEMCPY, OP_EVENT_RECORD, OP_STREAM_WAIT, OP_BN_FWD_TRAIN, OP_BN_FWD_EVAL, OP_BN_BWD, OP_BILINEAR_FWD, OP_BILINEAR_BWD,
 OP_HEAD_SOFTMAX_FWD, OP_HEAD_SOFTMAX_BWD, OP_BN_FINALIZE, OP_BN_APPLY, OP_BN_EVAL_PREPARE,
 OP_HEAD_SOFTMAX_FINALIZE, OP_BCAST_MULTIMEM) = range(32)


class Op(C.Structure):
    _fields_ = [("kind", C.c_int), ("stream", C.c_int), ("p", C.c_void_p * 12), ("i", C.c_longlong * 8),
                ("f", C.c_float * 6)]


_bound = False


def _bind():
    global _bound
    if _bound:
        return
    L = N.lib()
    vp, i32 = C.c_void_p, C.c_int
    L.dpt_program_create.argtypes = [i32, i32, C.POINTER(vp)]
    L.dpt_program_create.restype = i32
    L.dpt_program_destroy.argtypes = [vp]
    L.dpt_program_destroy.restype = None
    L.dpt_program_add.argtypes = [vp, C.POINTER(Op)]
    L.dpt_program_add.restype = i32
    L.dpt_program_size.argtypes = [vp]
    L.dpt_program_size.restype = i32
    L.dpt_program_stream_wait.argtypes = [vp, i32, vp]
    L.dpt_program_stream_wait.restype = i32
    L.dpt_program_run.argtypes = [vp, i32, i32, vp, vp]
    L.dpt_program_run.restype = i32
    L.dpt_run_parallel.argtypes = [i32, C.POINTER(vp), C.POINTER(i32), C.POINTER(i32), C.POINTER(vp), C.POINTER(vp)]
    L.dpt_run_parallel.restype = i32
    _bound = True


def _ptr(t) -> Optional[int]:
    if t is None:
        return None
    if isinstance(t, int):
        return t
    return t.data_ptr()


# DPT_DEBUG_SKIP_OPS="bgrad,wgrad": leave out every op whose label contains one of the tokens.  Results are then
# meaningless; this exists to measure what an op kind really costs inside the back-to-back step.
_DEBUG_SKIP = [t for t in os.environ.get("DPT_DEBUG_SKIP_OPS", "").split(",") if t]


class Program:
    def __init__(self, device: torch.device, n_events: int = 8):
        _bind()
        self.device = torch.device(device)
        self._h = C.c_void_p()
        self._keep: List[object] = []
        self.labels: List[str] = []
        self.kinds: List[int] = []
        with torch.cuda.device(self.device):
            N.check(N.lib().dpt_program_create(self.device.index or 0, n_events, C.byref(self._h)))

    def __len__(self) -> int:
        return int(N.lib().dpt_program_size(self._h))

    @property
    def handle(self) -> int:
        return self._h.value

    def keep(self, *objs) -> None:
        self._keep.extend(objs)

    def _add(self, kind: int, label: str, p: Sequence = (), i: Sequence[int] = (), f: Sequence[float] = (),
             stream: int = 0) -> int:
        if _DEBUG_SKIP and any(tok in label for tok in _DEBUG_SKIP):
            return len(self)  # ablation runs only (scripts/ablate_step.py): the op is left out of the program
        op = Op()
        op.kind, op.stream = kind, stream
        for k, v in enumerate(p):
            op.p[k] = _ptr(v)
        for k, v in enumerate(i):
            op.i[k] = int(v)
        for k, v in enumerate(f):
            op.f[k] = float(v)
        idx = N.lib().dpt_program_add(self._h, C.byref(op))
        if idx < 0:
            raise N.NativeError(N.lib().dpt_last_error().decode())
        self._keep.extend(x for x in p if x is not None and not isinstance(x, int))
        self.labels.append(label)
        self.kinds.append(kind)
        return idx

    # ---- one method per op kind --------------------------------------------------------------------------
    def conv(self, plan: N.ConvPlan, label="conv", stream=0):
        self.keep(plan)
        return self._add(OP_CONV, label, [plan.handle], stream=stream)

    def wgrad(self, plan: N.WgradPlan, label="wgrad", stream=0):
        self.keep(plan)
        return self._add(OP_WGRAD, label, [plan.handle], stream=stream)

    def pack_input(self, src: torch.Tensor, dst: torch.Tensor, label="pack_input"):
        """fp32 (already normalised) or uint8 (raw pixels, scaled by 1/255 here) NCHW -> bf16 NHWC padded to 16 channels."""
        b, c, h, w = src.shape
        u8 = src.dtype == torch.uint8
        assert u8 or src.dtype == torch.float32
        return self._add(OP_PACK_INPUT, label, [src, dst], [b, c, h, w, int(u8)], [1.0 / 255.0 if u8 else 1.0])

    def maxpool(self, x, y, label="maxpool"):
        b, h, w, c = x.shape
        return self._add(OP_MAXPOOL, label, [x, y], [b, h, w, c])

    def pool_bwd(self, act, d_pool, d_skip, dz, relu_mask=True, label="pool_bwd"):
        b, h, w, c = act.shape
        return self._add(OP_POOL_BWD, label, [act, d_pool, d_skip, dz], [b, h, w, c, int(relu_mask)])

    def channel_sum(self, x, out, scratch, scale=1.0, label="bias_grad", accumulate=False):
        c = x.shape[-1]
        return self._add(OP_CHANNEL_SUM, label, [x, out, scratch],
                         [x.numel() // c, c, scratch.numel() * scratch.element_size(), int(accumulate)], [scale])

    def prep_weights(self, table: torch.Tensor, n_entries: int, label="prep_weights", stream=0):
        return self._add(OP_PREP_WEIGHTS, label, [table], [n_entries], stream=stream)

    def head_fwd(self, x, w, b, target, target_dtype, prob, partial, label="head_fwd"):
        c = x.shape[-1]
        return self._add(OP_HEAD_FWD, label, [x, w, b, target, prob, partial], [x.numel() // c, target_dtype, c])

    def head_finalize(self, partials: Sequence, blocks: int, total_pixels: int, dice_weight: float, state, loss_log=None,
                      log_slot=0, label="head_finalize"):
        p = list(partials) + [None] * (8 - len(partials)) + [state, loss_log]
        return self._add(OP_HEAD_FINALIZE, label, p, [len(partials), blocks, log_slot, total_pixels], [0.0, dice_weight])

    def head_bwd(self, x, w, b, target, target_dtype, state, dx, partial_w, dw, db, gscale, dice_weight, relu_mask=True,
                 label="head_bwd", accumulate=False):
        c = x.shape[-1]
        return self._add(OP_HEAD_BWD, label, [x, w, b, target, state, dx, partial_w, dw, db],
                         [x.numel() // c, target_dtype, c, int(relu_mask), int(accumulate)], [gscale, dice_weight])

    def opt_tick(self, opt_state, label="opt_tick", stream=0):
        return self._add(OP_OPT_TICK, label, [opt_state], stream=stream)

    def adam(self, p, g, m, v, n, opt_state, eps, wd, gscale=1.0, zero_grad=False, label="adam", stream=0):
        return self._add(OP_ADAM, label, [p, g, m, v, opt_state], [n, int(zero_grad)], [eps, wd, gscale], stream=stream)

    def rmsprop(self, p, g, sq, buf, n, opt_state, alpha, eps, wd, momentum, gscale=1.0, zero_grad=False,
                label="rmsprop", stream=0):
        return self._add(OP_RMSPROP, label, [p, g, sq, buf, opt_state], [n, int(zero_grad)],
                         [eps, wd, gscale, alpha, momentum], stream=stream)

    def fused_allreduce(self, team_struct, desc_struct, label="fused_allreduce_opt", stream=0):
        self.keep(team_struct, desc_struct)
        return self._add(OP_FUSED_AR, label, [C.addressof(team_struct), C.addressof(desc_struct)], stream=stream)

    def bcast(self, team_struct, bufs_array, n, root=0, blocks=0, label="bcast", stream=0):
        self.keep(team_struct, bufs_array)
        return self._add(OP_BCAST, label, [C.addressof(team_struct), C.addressof(bufs_array)], [n, root, blocks],
                         stream=stream)

    def bcast_multimem(self, team_struct, root_src: int, mc_ptr: int, n, root=0, blocks=0, label="bcast (multimem.st)", stream=0):
        self.keep(team_struct)
        return self._add(OP_BCAST_MULTIMEM, label, [C.addressof(team_struct), root_src, mc_ptr], [n, root, blocks], stream=stream)

    def allreduce_sum(self, team_struct, bufs_array, n, scale=1.0, blocks=0, label="allreduce_sum", stream=0):
        self.keep(team_struct, bufs_array)
        return self._add(OP_ALLREDUCE_SUM, label, [C.addressof(team_struct), C.addressof(bufs_array)], [n, blocks],
                         [scale], stream=stream)

    def memset(self, t: torch.Tensor, value=0, label="memset", stream=0):
        return self._add(OP_MEMSET, label, [t], [t.numel() * t.element_size(), value], stream=stream)

    def memcpy(self, dst, src, nbytes, label="memcpy", stream=0):
        return self._add(OP_MEMCPY, label, [dst, src], [nbytes], stream=stream)

    # ---- north-star model family ------------------------------------------------------------------------
    def bn_fwd_train(self, y_raw, y, gamma, beta, running_mean, running_var, bn, scratch, eps=1e-5, momentum=0.1, relu=True,
                     label="bn+relu"):
        c = y_raw.shape[-1]
        return self._add(OP_BN_FWD_TRAIN, label, [y_raw, y, gamma, beta, running_mean, running_var, bn, None, scratch],
                         [y_raw.numel() // c, c, int(relu), scratch.numel() * scratch.element_size()], [eps, momentum])

    def bn_fwd_eval(self, y_raw, y, gamma, beta, running_mean, running_var, bn, eps=1e-5, relu=True, label="bn(eval)+relu"):
        c = y_raw.shape[-1]
        return self._add(OP_BN_FWD_EVAL, label, [y_raw, y, gamma, beta, running_mean, running_var, bn],
                         [y_raw.numel() // c, c, int(relu)], [eps])

    def bn_finalize(self, partial, rows_per_tile, nt, gamma, beta, running_mean, running_var, bn, pixels, eps=1e-5,
                    momentum=0.1, label="bn finalize"):
        """Batch statistics from the per-CTA rows the conv's statistics epilogue wrote (ConvPlan(stats=...)) -> bn [4][c]."""
        c = gamma.numel()
        return self._add(OP_BN_FINALIZE, label, [partial, gamma, beta, running_mean, running_var, bn],
                         [rows_per_tile, nt, pixels, c], [eps, momentum])

    def bn_apply(self, y_raw, y, bn, pooled=None, relu=True, label="bn apply+relu"):
        """y = relu(y_raw*scale+shift); `pooled` additionally receives MaxPool2d(2,2) of y from the same pass."""
        b, h, w, c = y_raw.shape
        return self._add(OP_BN_APPLY, label, [y_raw, y, pooled, bn], [b, h, w, c, int(relu)])

    def bn_eval_prepare(self, gamma, beta, running_mean, running_var, bn, eps=1e-5, label="bn eval scale/shift"):
        return self._add(OP_BN_EVAL_PREPARE, label, [gamma, beta, running_mean, running_var, bn], [gamma.numel()], [eps])

    def bn_bwd(self, y_raw, dy, dz, bn, dgamma, dbeta, scratch, accumulate=False, label="bn bwd"):
        c = y_raw.shape[-1]
        return self._add(OP_BN_BWD, label, [y_raw, dy, dz, bn, dgamma, dbeta, scratch],
                         [y_raw.numel() // c, c, int(accumulate), scratch.numel() * scratch.element_size()])

    def bilinear_fwd(self, x, out, label="bilinear x2"):
        b, h, w, c = x.shape
        return self._add(OP_BILINEAR_FWD, label, [x, out], [b, h, w, c])

    def bilinear_bwd(self, dout, din, label="bilinear x2 bwd"):
        b, h, w, c = din.shape
        return self._add(OP_BILINEAR_BWD, label, [dout, din], [b, h, w, c])

    def head_softmax_fwd(self, x, w, b, target, target_dtype, logits, partial, state, classes, label="head softmax+loss"):
        bsz, h, wd, c = x.shape
        return self._add(OP_HEAD_SOFTMAX_FWD, label, [x, w, b, target, logits, partial, state],
                         [bsz * h * wd, h * wd, target_dtype, c, classes])

    def head_softmax_finalize(self, partials: Sequence, total_pixels: int, state, label="softmax loss finalize"):
        p = list(partials) + [None] * (8 - len(partials)) + [state]
        return self._add(OP_HEAD_SOFTMAX_FINALIZE, label, p, [len(partials), total_pixels])

    def head_softmax_bwd(self, x, w, b, target, target_dtype, state, dx, dw, db, scratch, classes, gscale, accumulate=False,
                         label="head softmax bwd"):
        c = x.shape[-1]
        return self._add(OP_HEAD_SOFTMAX_BWD, label, [x, w, b, target, state, dx, dw, db, scratch],
                         [x.numel() // c, target_dtype, c, classes, int(accumulate), scratch.numel() * scratch.element_size()],
                         [gscale])

    def record(self, event: int, stream=0, label="record"):
        return self._add(OP_EVENT_RECORD, label, [], [event], stream=stream)

    def wait(self, event: int, stream=0, label="wait"):
        return self._add(OP_STREAM_WAIT, label, [], [event], stream=stream)

    def stream_wait(self, event: int, stream: torch.cuda.Stream) -> None:
        """`stream` waits for the program's event `event` as last recorded by a `record` op."""
        N.check(N.lib().dpt_program_stream_wait(self._h, event, stream.cuda_stream))

    # ---- execution ---------------------------------------------------------------------------------------
    def run(self, first: int = 0, last: int = -1, stream: Optional[torch.cuda.Stream] = None,
            aux_stream: Optional[torch.cuda.Stream] = None) -> None:
        # the stream of the PROGRAM's device (not of whatever device happens to be current)
        s = (stream if stream is not None else torch.cuda.current_stream(self.device)).cuda_stream
        a = aux_stream.cuda_stream if aux_stream is not None else s
        N.check(N.lib().dpt_program_run(self._h, first, last, s, a))

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value and N._lib is not None:
            N._lib.dpt_program_destroy(h)
            self._h = C.c_void_p()


def run_parallel(programs: Sequence[Program], streams: Sequence[torch.cuda.Stream],
                 aux_streams: Optional[Sequence[torch.cuda.Stream]] = None, first: Optional[Sequence[int]] = None,
                 last: Optional[Sequence[int]] = None) -> None:
    """Run one program per device concurrently from native worker threads (no GIL in the launch loops)."""
    _bind()
    n = len(programs)
    hs = (C.c_void_p * n)(*[p.handle for p in programs])
    sm = (C.c_void_p * n)(*[s.cuda_stream for s in streams])
    sa = (C.c_void_p * n)(*[(aux_streams[k] if aux_streams else streams[k]).cuda_stream for k in range(n)])
    fi = (C.c_int * n)(*(first or [0] * n))
    la = (C.c_int * n)(*(last or [-1] * n))
    N.check(N.lib().dpt_run_parallel(n, hs, fi, la, sm, sa))
