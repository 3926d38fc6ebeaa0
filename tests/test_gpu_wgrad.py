"""tcgen05 weight-gradient kernels (MN-major operands, split-K + deterministic reduce) vs PyTorch fp32."""
import pytest
import torch

from distributedpytorch_b200.ops import native as N
from distributedpytorch_b200.ops import reference as R
from gpu_util import assert_close, rand_bf16

pytestmark = pytest.mark.gpu


def _wgrad3x3(mode, B, H, W, cins, cout, ksplit=0):
    xs = [rand_bf16(B, H, W, c, seed=61 + i) for i, c in enumerate(cins)]
    dz = rand_bf16(B, H, W, cout, scale=0.5, seed=65)
    cin = sum(cins)
    dw = torch.full((cout, cin, 3, 3), float("nan"), device="cuda")
    db = torch.full((cout,), float("nan"), device="cuda")  # fused bias gradient: per-channel sum of dz
    plan = N.WgradPlan(mode=mode, batch=B, h=H, w=W, z=dz, xs=xs, dw=dw, strides=(cin * 9, 9, 1), ksplit=ksplit, dbias=db)
    plan.launch()
    torch.cuda.synchronize()
    ref = R.conv3x3_wgrad(torch.cat(xs, dim=-1), dz)
    scale = ref.abs().max().item()
    assert_close(dw, ref, 2e-3 * scale + 1e-3, 1e-2, f"wgrad mode={mode} {cins}->{cout} {B}x{H}x{W}")
    ref_b = dz.float().sum(dim=(0, 1, 2))
    assert_close(db, ref_b, 1e-3 * ref_b.abs().max().item() + 1e-2, 1e-3, f"fused bgrad mode={mode} {cins}->{cout}")
    # bit-reproducible
    dw2 = torch.empty_like(dw)
    db2 = torch.empty_like(db)
    plan2 = N.WgradPlan(mode=mode, batch=B, h=H, w=W, z=dz, xs=xs, dw=dw2, strides=(cin * 9, 9, 1), ksplit=ksplit, dbias=db2)
    plan2.launch()
    torch.cuda.synchronize()
    assert torch.equal(dw, dw2), "wgrad is not deterministic"
    assert torch.equal(db, db2), "fused bias gradient is not deterministic"
    # without the bias output the weight gradient is unchanged
    dw3 = torch.empty_like(dw)
    N.WgradPlan(mode=mode, batch=B, h=H, w=W, z=dz, xs=xs, dw=dw3, strides=(cin * 9, 9, 1), ksplit=ksplit).launch()
    torch.cuda.synchronize()
    assert torch.equal(dw, dw3)


@pytest.mark.parametrize("cins,cout", [((32,), 32), ((64,), 64), ((32,), 128)])
def test_wgrad_tapbox(cins, cout):
    _wgrad3x3(N.CONV_3X3_TAPBOX, 2, 40, 60, cins, cout)


@pytest.mark.parametrize("cins,cout", [
    ((32,), 32), ((64,), 64), ((64,), 128), ((128,), 256), ((16,), 32), ((32, 32), 32), ((64, 64), 64),
    ((256,), 512),
])
def test_wgrad_halo(cins, cout):
    _wgrad3x3(N.CONV_3X3, 2, 40, 60, cins, cout)


def test_wgrad_full_res():
    _wgrad3x3(N.CONV_3X3, 2, 160, 240, (32,), 32)


@pytest.mark.parametrize("cin,cout", [(64, 32), (256, 128)])
def test_wgrad_convT(cin, cout):
    B, H, W = 2, 24, 40
    x = rand_bf16(B, H, W, cin, seed=71)
    dy = rand_bf16(B, 2 * H, 2 * W, cout, scale=0.5, seed=72)
    dw = torch.full((cin, cout, 2, 2), float("nan"), device="cuda")
    db = torch.full((cout,), float("nan"), device="cuda")
    plan = N.WgradPlan(mode=N.CONV_2X2S2, batch=B, h=H, w=W, z=x, xs=[dy], dw=dw, strides=(cout * 4, 4, 1), dbias=db)
    plan.launch()
    torch.cuda.synchronize()
    ref = R.convT2x2_wgrad(x, dy)
    scale = ref.abs().max().item()
    assert_close(dw, ref, 2e-3 * scale + 1e-3, 1e-2, f"convT wgrad {cin}->{cout}")
    ref_b = dy.float().sum(dim=(0, 1, 2))  # fused bias gradient = per-channel sum of the output gradient
    assert_close(db, ref_b, 1e-3 * ref_b.abs().max().item() + 1e-2, 1e-3, f"convT fused bgrad {cin}->{cout}")
