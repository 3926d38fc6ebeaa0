"""Kernels of the north-star model family (BatchNorm+ReLU fwd/bwd, bilinear x2, softmax CE+Dice head, conv without bias
at 64..1024 channels) against PyTorch fp32 (same op)."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from distributedpytorch_b200.ops import native as N
from distributedpytorch_b200.ops import reference as R
from distributedpytorch_b200.utils.loss import CEDiceLoss
from gpu_util import assert_close, rand_bf16, rand_f32

pytestmark = pytest.mark.gpu
vp, i32, i64, f32, sz = C.c_void_p, C.c_int, C.c_longlong, C.c_float, C.c_size_t


def _lib():
    L = N.lib()
    if getattr(L, "_mil_bound", False):
        return L
    L.dpt_bn_scratch_bytes.argtypes = [i32]
    L.dpt_bn_scratch_bytes.restype = sz
    L.dpt_bn_forward_train.argtypes = [vp, vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, f32, i32, vp, sz, vp]
    L.dpt_bn_forward_eval.argtypes = [vp, vp, vp, vp, vp, vp, vp, i64, i32, f32, i32, vp]
    L.dpt_bn_backward.argtypes = [vp, vp, vp, vp, vp, vp, i64, i32, i32, vp, sz, vp]
    L.dpt_bilinear_up2x_fwd.argtypes = [vp, vp, i32, i32, i32, i32, vp]
    L.dpt_bilinear_up2x_bwd.argtypes = [vp, vp, i32, i32, i32, i32, vp]
    L.dpt_head_softmax_fwd.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, i64, i64, i32, i32, vp]
    L.dpt_head_softmax_scratch_bytes.argtypes = [i64, i32, i32]
    L.dpt_head_softmax_scratch_bytes.restype = sz
    L.dpt_head_softmax_bwd.argtypes = [vp, vp, vp, vp, i32, vp, vp, vp, vp, i64, i32, i32, f32, i32, vp, sz, vp]
    L._mil_bound = True
    return L


@pytest.mark.parametrize("c", [64, 256])
def test_batchnorm_relu_forward_backward(c):
    L = _lib()
    B, H, W = 3, 24, 40
    P = B * H * W
    y_raw = rand_bf16(B, H, W, c, scale=2.0, seed=1)
    gamma = rand_f32(c, seed=2) * 0.5 + 1.0
    beta = rand_f32(c, seed=3) * 0.2
    rm, rv = torch.zeros(c, device="cuda"), torch.ones(c, device="cuda")
    bn = torch.empty(4, c, device="cuda")
    y = torch.empty_like(y_raw)
    nbytes = int(L.dpt_bn_scratch_bytes(c)) + 2 * c * 4
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    s = N.stream_ptr()
    N.check(L.dpt_bn_forward_train(y_raw.data_ptr(), y.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                   bn.data_ptr(), None, P, c, 1e-5, 0.1, 1, scratch.data_ptr(), nbytes, s))
    dy = rand_bf16(B, H, W, c, seed=4)
    dz = torch.empty_like(y_raw)
    dgamma, dbeta = torch.empty(c, device="cuda"), torch.empty(c, device="cuda")
    N.check(L.dpt_bn_backward(y_raw.data_ptr(), dy.data_ptr(), dz.data_ptr(), bn.data_ptr(), dgamma.data_ptr(), dbeta.data_ptr(), P,
                              c, 0, scratch.data_ptr(), nbytes, s))
    torch.cuda.synchronize()
    # oracle
    m = torch.nn.BatchNorm2d(c).cuda()
    with torch.no_grad():
        m.weight.copy_(gamma)
        m.bias.copy_(beta)
    x = R.nhwc_to_nchw(y_raw.float()).requires_grad_(True)
    out = F.relu(m(x))
    out.backward(R.nhwc_to_nchw(dy.float()))
    assert_close(y, R.nchw_to_nhwc(out.detach()), 2e-2, 2e-2, "bn+relu forward")
    assert_close(rm, m.running_mean, 1e-4, 1e-3, "running_mean")
    assert_close(rv, m.running_var, 1e-4, 1e-3, "running_var")
    assert_close(dgamma, m.weight.grad, 5e-2, 1e-2, "dgamma")
    assert_close(dbeta, m.bias.grad, 5e-2, 1e-2, "dbeta")
    assert_close(dz, R.nchw_to_nhwc(x.grad), 2e-2, 3e-2, "bn backward dz")
    # eval mode
    y2 = torch.empty_like(y_raw)
    N.check(L.dpt_bn_forward_eval(y_raw.data_ptr(), y2.data_ptr(), gamma.data_ptr(), beta.data_ptr(), rm.data_ptr(), rv.data_ptr(),
                                  bn.data_ptr(), P, c, 1e-5, 1, s))
    torch.cuda.synchronize()
    m.eval()
    with torch.no_grad():
        assert_close(y2, R.nchw_to_nhwc(F.relu(m(x.detach()))), 2e-2, 2e-2, "bn eval")


@pytest.mark.parametrize("c,relu", [(64, 1), (128, 0), (1024, 1)])
def test_bn_apply_with_fused_maxpool(c, relu):
    """dpt_bn_apply: y = relu(x*scale+shift) and, in the same pass, MaxPool2d(2,2) of y."""
    B, H, W = 2, 12, 20
    x = rand_bf16(B, H, W, c, scale=2.0, seed=21)
    bn = torch.zeros(4, c, device="cuda")
    bn[2] = rand_f32(c, seed=22) * 0.5 + 1.0
    bn[3] = rand_f32(c, seed=23) * 0.3
    y = torch.empty_like(x)
    pooled = torch.empty(B, H // 2, W // 2, c, dtype=torch.bfloat16, device="cuda")
    N.check(N.lib().dpt_bn_apply(x.data_ptr(), y.data_ptr(), pooled.data_ptr(), bn.data_ptr(), B, H, W, c, relu, N.stream_ptr()))
    y2 = torch.empty_like(x)
    N.check(N.lib().dpt_bn_apply(x.data_ptr(), y2.data_ptr(), None, bn.data_ptr(), B, H, W, c, relu, N.stream_ptr()))
    torch.cuda.synchronize()
    ref = x.float() * bn[2] + bn[3]
    if relu:
        ref = ref.clamp_min(0)
    assert_close(y, ref, 1e-2, 1e-2, "bn apply")
    assert torch.equal(y, y2)
    ref_pool = F.max_pool2d(R.nhwc_to_nchw(y.float()), 2)
    assert torch.equal(pooled.float(), R.nchw_to_nhwc(ref_pool)), "fused max-pool differs from pooling the stored activation"


@pytest.mark.parametrize("h,w,c", [(12, 20, 64), (5, 7, 128)])
def test_bilinear_up2x(h, w, c):
    L = _lib()
    B = 2
    x = rand_bf16(B, h, w, c, seed=5)
    out = torch.empty(B, 2 * h, 2 * w, c, dtype=torch.bfloat16, device="cuda")
    s = N.stream_ptr()
    N.check(L.dpt_bilinear_up2x_fwd(x.data_ptr(), out.data_ptr(), B, h, w, c, s))
    dout = rand_bf16(B, 2 * h, 2 * w, c, seed=6)
    din = torch.empty_like(x)
    N.check(L.dpt_bilinear_up2x_bwd(dout.data_ptr(), din.data_ptr(), B, h, w, c, s))
    torch.cuda.synchronize()
    xr = R.nhwc_to_nchw(x.float()).requires_grad_(True)
    ref = F.interpolate(xr, scale_factor=2, mode="bilinear", align_corners=True)
    ref.backward(R.nhwc_to_nchw(dout.float()))
    assert_close(out, R.nchw_to_nhwc(ref.detach()), 1e-2, 1e-2, "bilinear fwd")
    assert_close(din, R.nchw_to_nhwc(xr.grad), 3e-2, 2e-2, "bilinear bwd")


@pytest.mark.parametrize("classes", [2, 3])
def test_softmax_ce_dice_head(classes):
    L = _lib()
    B, H, W, c = 2, 24, 40, 64
    P = B * H * W
    x = torch.relu(rand_bf16(B, H, W, c, seed=7).float()).to(torch.bfloat16)
    w = rand_f32(classes, c, scale=0.2, seed=8)
    b = rand_f32(classes, scale=0.1, seed=9)
    t = torch.randint(0, classes, (B, H, W), device="cuda", generator=torch.Generator(device="cuda").manual_seed(10))
    tgt = t.to(torch.uint8)
    logits = torch.empty(B, classes, H, W, device="cuda")
    blocks = L.dpt_head_blocks()
    partial = torch.empty(blocks, 4, device="cuda")
    state = torch.zeros(8, device="cuda")
    s = N.stream_ptr()
    N.check(L.dpt_head_softmax_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), tgt.data_ptr(), 1, logits.data_ptr(), partial.data_ptr(),
                                   state.data_ptr(), P, H * W, c, classes, s))
    nbytes = int(L.dpt_head_softmax_scratch_bytes(P, c, classes))
    scratch = torch.empty(nbytes, dtype=torch.uint8, device="cuda")
    dx = torch.empty_like(x)
    dw, db = torch.empty(classes, c, device="cuda"), torch.empty(classes, device="cuda")
    N.check(L.dpt_head_softmax_bwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), tgt.data_ptr(), 1, state.data_ptr(), dx.data_ptr(),
                                   dw.data_ptr(), db.data_ptr(), P, c, classes, 3.0, 0, scratch.data_ptr(), nbytes, s))
    torch.cuda.synchronize()
    xr = x.float().requires_grad_(True)
    wr, br = w.clone().requires_grad_(True), b.clone().requires_grad_(True)
    lg = (xr @ wr.t() + br).permute(0, 3, 1, 2)
    loss = CEDiceLoss(classes)(lg, t)
    (3.0 * loss).backward()
    assert_close(logits, lg.detach(), 1e-4, 1e-4, "logits")
    assert_close(state[4], loss.detach(), 1e-4, 1e-4, "ce+dice loss")
    assert_close(dw, wr.grad, 1e-4, 3e-3, "outc dw")
    assert_close(db, br.grad, 1e-4, 3e-3, "outc db")
    assert_close(dx, xr.grad, 1e-6, 2e-2, "outc dx")


def test_wide_biasless_conv_like_down4():
    # 512 -> 1024 at the bottom of the milesial UNet: N tiling over 8 tiles, 8 K chunks, no bias, no ReLU
    B, H, W = 1, 40, 60
    x = rand_bf16(B, H, W, 512, seed=11)
    w = rand_f32(1024, 512, 3, 3, scale=(9 * 512) ** -0.5 * 2, seed=12).to(torch.bfloat16).float()
    out = torch.empty(B, H, W, 1024, dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[x], weight=R.conv3x3_weight_fwd(w), cout=1024, dsts=[out])
    plan.launch()
    torch.cuda.synchronize()
    assert_close(out, R.conv3x3([x], w, None, False), 2e-2, 2e-2, f"512->1024 conv {plan.info()}")
