"""Native engine for the north-star model family (BatchNorm UNet, 2 classes, ConvTranspose or bilinear) vs the same
module on fp32 torch ops + autograd."""
import copy

import pytest
import torch

from distributedpytorch_b200.models import UNet
from distributedpytorch_b200.utils.loss import CEDiceLoss

pytestmark = pytest.mark.gpu


def _cos(a, b):
    a, b = a.flatten().float(), b.flatten().float()
    return float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-30))


@pytest.mark.parametrize("bilinear", [False, True])
def test_milesial_engine_matches_torch(bilinear):
    from distributedpytorch_b200.engine.milesial import NativeMilesialEngine

    torch.manual_seed(0)
    B, H, W = 4, 64, 96
    ref = UNet(variant="milesial", n_classes=2, bilinear=bilinear).cuda()
    model = copy.deepcopy(ref)
    eng = NativeMilesialEngine(model, B, H, W, "cuda", lr=1e-3, optimizer="adam", loss_scale=1.0)
    g = torch.Generator(device="cuda").manual_seed(1)
    x = torch.rand(B, 3, H, W, device="cuda", generator=g)
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    t = (((yy - H / 2) ** 2 / (H / 3) ** 2 + (xx - W / 2) ** 2 / (W / 3) ** 2) < 1).long().expand(B, H, W).contiguous()
    ref.train()
    loss_ref = CEDiceLoss(2)(ref(x), t)
    loss_ref.backward()
    loss = eng.forward_backward(x, t.to(torch.uint8))
    torch.cuda.synchronize()
    assert abs(loss.item() - loss_ref.item()) < 0.03 * abs(loss_ref.item()) + 5e-3, (loss.item(), loss_ref.item())
    report, bad = [], []
    for (n, p_ref), (_, p) in zip(ref.named_parameters(), model.named_parameters()):
        c = _cos(p.grad, p_ref.grad)
        ratio = p.grad.norm().item() / (p_ref.grad.norm().item() + 1e-30)
        report.append(f"{n:50s} cos {c:.4f} ratio {ratio:.3f} |g| {p_ref.grad.norm().item():.3e}")
        # bf16 activations/gradients through 18 BatchNorm layers at random init: norms agree to ~1%, directions drift
        # smoothly with depth (0.99+ next to the loss, ~0.88 at the first layer); a structural error breaks the norms
        if not (c > 0.85 and 0.85 < ratio < 1.18):
            bad.append(n)
    assert not bad, "\n".join(report)
    # running statistics were updated like torch's
    for (n, b_ref), (_, b) in zip(ref.named_buffers(), model.named_buffers()):
        if "running" in n:
            assert torch.allclose(b, b_ref, atol=2e-2, rtol=5e-2), n
    # a few optimizer steps reduce the loss; eval-mode forward agrees with torch eval on the updated weights
    losses = [eng.train_step(x, t.to(torch.uint8)).item() for _ in range(6)]
    assert losses[-1] < losses[0], losses
    model.eval()
    with torch.no_grad():
        lg_ref = model(x)  # the module's parameters ARE the engine's arena
    lg = eng.forward(x)
    torch.cuda.synchronize()
    assert (lg - lg_ref).abs().max().item() < 0.15 * lg_ref.abs().max().item() + 0.05
