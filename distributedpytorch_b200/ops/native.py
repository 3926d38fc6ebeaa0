"""ctypes binding of libdpt_b200.so (the sm_100a kernels; C ABI in include/dpt_b200.h).

The library is built in-tree by `make` (see __graft_entry__.build) into distributedpytorch_b200/_native/.
On a machine with a GPU the ops in this package REQUIRE it: `lib()` raises if the shared object is
missing instead of silently falling back to PyTorch kernels.
"""
from __future__ import annotations

import ctypes as C
import os
from typing import Optional, Sequence

import torch

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(os.path.dirname(_HERE), "_native", "libdpt_b200.so")

# enums (keep in sync with include/dpt_b200.h)
CONV_3X3, CONV_3X3_TAPBOX, CONV_2X2S2, CONV_1X1 = 0, 1, 2, 3
OUT_NHWC, OUT_SHUFFLE2 = 0, 1
OPT_ADAM, OPT_RMSPROP = 0, 1
AR_TWO_SHOT, AR_ONE_SHOT, AR_MULTIMEM = 0, 1, 2
MAX_PEERS = 8
COMM_MAX_BLOCKS = 128
PAD_BYTES = COMM_MAX_BLOCKS * MAX_PEERS * 4
EPOCH_BYTES = COMM_MAX_BLOCKS * 4


class ConvDesc(C.Structure):
    _fields_ = [
        ("mode", C.c_int), ("out_mode", C.c_int),
        ("batch", C.c_int), ("h", C.c_int), ("w", C.c_int),
        ("num_src", C.c_int),
        ("src", C.c_void_p * 2), ("src_c", C.c_int * 2),
        ("weight", C.c_void_p), ("cout", C.c_int),
        ("bias", C.c_void_p), ("relu", C.c_int),
        ("num_dst", C.c_int),
        ("dst", C.c_void_p * 2), ("dst_c", C.c_int * 2), ("mask", C.c_void_p * 2),
        ("stats", C.c_void_p), ("pool_dst", C.c_void_p),
        ("tile_s", C.c_int), ("tile_n", C.c_int),
        ("src_h", C.c_int * 2), ("src_w", C.c_int * 2), ("src_y0", C.c_int * 2), ("src_x0", C.c_int * 2),
        ("cluster", C.c_int),
    ]


class WgradDesc(C.Structure):
    _fields_ = [
        ("mode", C.c_int), ("batch", C.c_int), ("h", C.c_int), ("w", C.c_int),
        ("z", C.c_void_p), ("z_c", C.c_int),
        ("num_x", C.c_int), ("x", C.c_void_p * 2), ("x_c", C.c_int * 2),
        ("dw", C.c_void_p),
        ("dw_stride_m", C.c_longlong), ("dw_stride_n", C.c_longlong), ("dw_stride_t", C.c_longlong),
        ("scale", C.c_float),
        ("scratch", C.c_void_p), ("scratch_bytes", C.c_size_t),
        ("ksplit", C.c_int), ("n_valid", C.c_int), ("accumulate", C.c_int),
        ("dbias", C.c_void_p),
    ]


class PrepEntry(C.Structure):
    _fields_ = [
        ("w", C.c_void_p), ("bias", C.c_void_p), ("wf", C.c_void_p), ("wd", C.c_void_p), ("bias4", C.c_void_p),
        ("cout", C.c_int), ("cin", C.c_int), ("cin_pad", C.c_int), ("kind", C.c_int), ("oscale", C.c_void_p),
    ]


class Team(C.Structure):
    _fields_ = [("rank", C.c_int), ("world", C.c_int), ("pads", C.c_void_p * MAX_PEERS), ("epoch", C.c_void_p)]


class FusedArDesc(C.Structure):
    _fields_ = [
        ("grads", C.c_void_p * MAX_PEERS), ("params", C.c_void_p * MAX_PEERS),
        ("mc_grads", C.c_void_p), ("mc_params", C.c_void_p),
        ("state0", C.c_void_p), ("state1", C.c_void_p), ("opt_state", C.c_void_p),
        ("n", C.c_longlong),
        ("algo", C.c_int), ("optimizer", C.c_int), ("blocks", C.c_int),
        ("eps", C.c_float), ("weight_decay", C.c_float), ("grad_scale", C.c_float),
        ("alpha", C.c_float), ("momentum", C.c_float),
    ]


_lib: Optional[C.CDLL] = None


class NativeError(RuntimeError):
    pass


def available() -> bool:
    return os.path.exists(LIB_PATH)


def lib() -> C.CDLL:
    """Load (once) and return the native library; raises if it has not been built."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise NativeError(
            f"{LIB_PATH} not found: run `make -j` (or `python -c 'import __graft_entry__ as g; g.build()'`) "
            "at the repository root. The B200 kernels are mandatory on GPU; there is no silent fallback.")
    L = C.CDLL(LIB_PATH)
    vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t
    L.dpt_last_error.restype = C.c_char_p
    L.dpt_version.restype = i32
    L.dpt_launch_count.restype = i64

    def sig(name, args, res=i32):
        fn = getattr(L, name)
        fn.argtypes = args
        fn.restype = res

    sig("dpt_conv_plan_create", [C.POINTER(ConvDesc), C.POINTER(vp)])
    sig("dpt_conv_plan_launch", [vp, vp])
    sig("dpt_conv_plan_destroy", [vp], None)
    sig("dpt_conv_plan_info", [vp, C.POINTER(i32 * 9)])
    sig("dpt_conv_plan_cluster", [vp])
    sig("dpt_wgrad_scratch_bytes", [C.POINTER(WgradDesc)], sz)
    sig("dpt_wgrad_plan_create", [C.POINTER(WgradDesc), C.POINTER(vp)])
    sig("dpt_wgrad_plan_launch", [vp, vp])
    sig("dpt_wgrad_plan_destroy", [vp], None)
    sig("dpt_pack_input", [vp, vp, i32, i32, i32, i32, vp])
    sig("dpt_pack_input_ex", [vp, i32, C.c_float, vp, i32, i32, i32, i32, vp])
    sig("dpt_maxpool2x2", [vp, vp, i32, i32, i32, i32, vp])
    sig("dpt_pool_bwd_fused", [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp])
    sig("dpt_channel_sum_scratch_bytes", [i32], sz)
    sig("dpt_channel_sum", [vp, vp, i64, i32, f32, vp, sz, vp])
    sig("dpt_prep_weights", [vp, i32, vp])
    sig("dpt_head_blocks", [], i32)
    sig("dpt_head_sigmoid_fwd", [vp, vp, vp, vp, i32, vp, vp, i64, i32, vp])
    sig("dpt_head_finalize", [C.POINTER(vp), C.POINTER(i32), i32, C.c_double, f32, vp, vp, i32, vp])
    sig("dpt_head_sigmoid_bwd", [vp, vp, vp, vp, i32, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp])
    sig("dpt_opt_init", [vp, f32, i32, f32, f32, vp])
    sig("dpt_opt_set_lr", [vp, f32, vp])
    sig("dpt_opt_tick", [vp, vp])
    sig("dpt_adam_step", [vp, vp, vp, vp, i64, vp, f32, f32, f32, i32, vp])
    sig("dpt_rmsprop_step", [vp, vp, vp, vp, i64, vp, f32, f32, f32, f32, f32, i32, vp])
    sig("dpt_fused_allreduce_opt", [C.POINTER(Team), C.POINTER(FusedArDesc), vp])
    sig("dpt_team_bcast", [C.POINTER(Team), C.POINTER(vp), i64, i32, i32, vp])
    sig("dpt_team_allreduce_sum", [C.POINTER(Team), C.POINTER(vp), i64, f32, i32, vp])
    sig("dpt_team_bcast_multimem", [C.POINTER(Team), vp, vp, i64, i32, i32, vp])
    sig("dpt_enable_peer_access", [i32, i32])
    sig("dpt_bn_scratch_bytes", [i32], sz)
    sig("dpt_bn_finalize", [vp, i32, i32, vp, vp, vp, vp, vp, i64, i32, f32, f32, vp])
    sig("dpt_bn_apply", [vp, vp, vp, vp, i32, i32, i32, i32, i32, vp])
    sig("dpt_bn_eval_prepare", [vp, vp, vp, vp, vp, i32, f32, vp])
    sig("dpt_head_softmax_scratch_bytes", [i64, i32, i32], sz)
    sig("dpt_set_sm_reserve", [i32])
    sig("dpt_get_sm_count", [])
    sig("dpt_head_softmax_finalize", [C.POINTER(vp), i32, C.c_double, vp, vp])
    _lib = L
    return L


def check(rc: int) -> None:
    if rc != 0:
        raise NativeError(lib().dpt_last_error().decode("utf-8", "replace"))


def launch_count() -> int:
    return int(lib().dpt_launch_count()) if _lib is not None or available() else 0


def ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


def stream_ptr(stream: Optional[torch.cuda.Stream] = None) -> int:
    s = stream if stream is not None else torch.cuda.current_stream()
    return s.cuda_stream


class ConvPlan:
    """Owns a dpt_conv_plan (tensor maps + launch geometry) and keeps the tensors it points at alive."""

    def __init__(self, *, mode: int, batch: int, h: int, w: int, srcs: Sequence[torch.Tensor], weight: torch.Tensor,
                 cout: int, dsts: Sequence[torch.Tensor], bias: Optional[torch.Tensor] = None, relu: bool = False,
                 masks: Sequence[Optional[torch.Tensor]] = (None, None), out_mode: int = OUT_NHWC,
                 stats: Optional[torch.Tensor] = None, tile_s: int = 0, tile_n: int = 0, cluster: int = 0,
                 src_offsets: Optional[Sequence[Optional[Sequence[int]]]] = None, pooled: Optional[torch.Tensor] = None):
        """`src_offsets[i] = (y0, x0)`: source i is larger than [batch, h, w] and the GEMM reads the window at that corner
        (center crop of a skip tensor as a view)."""
        d = ConvDesc()
        d.mode, d.out_mode = mode, out_mode
        d.batch, d.h, d.w = batch, h, w
        d.num_src = len(srcs)
        for i, s in enumerate(srcs):
            assert s.dtype == torch.bfloat16 and s.is_contiguous()
            d.src[i] = s.data_ptr()
            d.src_c[i] = s.shape[-1]
            off = src_offsets[i] if src_offsets is not None and i < len(src_offsets) else None
            if off is not None:
                d.src_h[i], d.src_w[i], d.src_y0[i], d.src_x0[i] = s.shape[1], s.shape[2], off[0], off[1]
        assert weight.dtype == torch.bfloat16 and weight.is_contiguous()
        d.weight, d.cout = weight.data_ptr(), cout
        d.bias = ptr(bias)
        d.relu = int(relu)
        d.num_dst = len(dsts)
        for i, t in enumerate(dsts):
            assert t.dtype == torch.bfloat16 and t.is_contiguous()
            d.dst[i] = t.data_ptr()
            d.dst_c[i] = t.shape[-1]
            m = masks[i] if i < len(masks) else None
            d.mask[i] = ptr(m)
        d.stats = ptr(stats)
        d.pool_dst = ptr(pooled)
        d.tile_s, d.tile_n, d.cluster = tile_s, tile_n, cluster
        self._keep = (list(srcs), weight, list(dsts), bias, list(masks), stats, pooled)
        self._h = C.c_void_p()
        check(lib().dpt_conv_plan_create(C.byref(d), C.byref(self._h)))

    def launch(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        check(lib().dpt_conv_plan_launch(self._h, stream_ptr(stream)))

    def info(self) -> dict:
        arr = (C.c_int * 9)()
        check(lib().dpt_conv_plan_info(self._h, C.byref(arr)))
        keys = ["grid", "smem", "tile_s", "tile_n", "a_stages", "b_stages", "b_resident", "acc_bufs", "chunk_ch"]
        out = dict(zip(keys, list(arr)))
        out["cluster"] = int(lib().dpt_conv_plan_cluster(self._h))
        return out

    @property
    def handle(self) -> int:
        return self._h.value

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value and _lib is not None:
            _lib.dpt_conv_plan_destroy(h)
            self._h = C.c_void_p()


class WgradPlan:
    def __init__(self, *, mode: int, batch: int, h: int, w: int, z: torch.Tensor, xs: Sequence[torch.Tensor],
                 dw: torch.Tensor, strides: Sequence[int], scratch: Optional[torch.Tensor] = None, scale: float = 1.0,
                 ksplit: int = 0, n_valid: int = 0, accumulate: bool = False, dbias: Optional[torch.Tensor] = None):
        d = WgradDesc()
        d.mode, d.batch, d.h, d.w = mode, batch, h, w
        assert z.dtype == torch.bfloat16 and z.is_contiguous()
        d.z, d.z_c = z.data_ptr(), z.shape[-1]
        d.num_x = len(xs)
        for i, x in enumerate(xs):
            assert x.dtype == torch.bfloat16 and x.is_contiguous()
            d.x[i] = x.data_ptr()
            d.x_c[i] = x.shape[-1]
        assert dw.dtype == torch.float32
        d.dw = dw.data_ptr()
        d.dw_stride_m, d.dw_stride_n, d.dw_stride_t = strides
        d.scale = scale
        d.ksplit = ksplit
        d.n_valid = n_valid
        d.accumulate = int(accumulate)
        if dbias is not None:  # bias gradient (column sums of the output gradient: z, or x for ConvTranspose) from the same kernels
            assert dbias.dtype == torch.float32 and dbias.numel() == (xs[0].shape[-1] if mode == CONV_2X2S2 else z.shape[-1])
            d.dbias = dbias.data_ptr()
        need = int(lib().dpt_wgrad_scratch_bytes(C.byref(d)))
        if scratch is None:
            scratch = torch.empty(need, dtype=torch.uint8, device=z.device)
        assert scratch.numel() * scratch.element_size() >= need, "wgrad scratch too small"
        d.scratch, d.scratch_bytes = scratch.data_ptr(), scratch.numel() * scratch.element_size()
        self.scratch_bytes = need
        self._keep = (z, list(xs), dw, scratch, dbias)
        self._h = C.c_void_p()
        check(lib().dpt_wgrad_plan_create(C.byref(d), C.byref(self._h)))

    @staticmethod
    def scratch_bytes_for(*, mode: int, batch: int, h: int, w: int, z_c: int, x_cs: Sequence[int], ksplit: int = 0) -> int:
        d = WgradDesc()
        d.mode, d.batch, d.h, d.w, d.z_c = mode, batch, h, w, z_c
        d.num_x = len(x_cs)
        for i, c in enumerate(x_cs):
            d.x_c[i] = c
        d.ksplit = ksplit
        return int(lib().dpt_wgrad_scratch_bytes(C.byref(d)))

    def launch(self, stream: Optional[torch.cuda.Stream] = None) -> None:
        check(lib().dpt_wgrad_plan_launch(self._h, stream_ptr(stream)))

    @property
    def handle(self) -> int:
        return self._h.value

    def __del__(self):
        h = getattr(self, "_h", None)
        if h is not None and h.value and _lib is not None:
            _lib.dpt_wgrad_plan_destroy(h)
            self._h = C.c_void_p()


# ---- thin functional wrappers (used by tests and by the Python-side executor) -------------------------

def pack_input(src: torch.Tensor, dst: torch.Tensor, stream=None) -> None:
    b, c, h, w = src.shape
    assert src.dtype in (torch.float32, torch.uint8) and src.is_contiguous() and dst.dtype == torch.bfloat16
    u8 = src.dtype == torch.uint8
    check(lib().dpt_pack_input_ex(src.data_ptr(), int(u8), 1.0 / 255.0 if u8 else 1.0, dst.data_ptr(), b, c, h, w,
                                  stream_ptr(stream)))


def maxpool2x2(x: torch.Tensor, y: torch.Tensor, stream=None) -> None:
    b, h, w, c = x.shape
    check(lib().dpt_maxpool2x2(x.data_ptr(), y.data_ptr(), b, h, w, c, stream_ptr(stream)))


def pool_bwd_fused(act, d_pool, d_skip, dz, relu_mask=True, stream=None) -> None:
    b, h, w, c = act.shape
    check(lib().dpt_pool_bwd_fused(act.data_ptr(), d_pool.data_ptr(), ptr(d_skip), dz.data_ptr(), b, h, w, c,
                                   int(relu_mask), stream_ptr(stream)))


def channel_sum(x: torch.Tensor, out: torch.Tensor, scale: float = 1.0, scratch: Optional[torch.Tensor] = None,
                stream=None) -> None:
    c = x.shape[-1]
    pixels = x.numel() // c
    need = int(lib().dpt_channel_sum_scratch_bytes(c))
    if scratch is None:
        scratch = torch.empty(need, dtype=torch.uint8, device=x.device)
    check(lib().dpt_channel_sum(x.data_ptr(), out.data_ptr(), pixels, c, scale, scratch.data_ptr(),
                                scratch.numel() * scratch.element_size(), stream_ptr(stream)))


def prep_weights(entries: Sequence[PrepEntry], device) -> torch.Tensor:
    """Upload a table of PrepEntry to the device; returns the uint8 tensor holding it."""
    arr = (PrepEntry * len(entries))(*entries)
    raw = bytes(arr)
    host = torch.frombuffer(bytearray(raw), dtype=torch.uint8)
    return host.to(device)


def run_prep_weights(table: torch.Tensor, n_entries: int, stream=None) -> None:
    check(lib().dpt_prep_weights(table.data_ptr(), n_entries, stream_ptr(stream)))
