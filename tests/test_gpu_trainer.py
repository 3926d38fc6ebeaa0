"""Trainer-level behaviour on the native engines: the loss curve records the mean of the last ten DIFFERENT losses
(utils/train_utils.py:75-79), the incomplete last batch is trained on (utils/train_utils.py:40), engines follow their own
device whatever device is current."""
import numpy as np
import pytest
import torch
from torch.utils.data import DataLoader

from distributedpytorch_b200.models import UNet
import importlib

F = importlib.import_module("distributedpytorch_b200.trainer.fit")  # the MODULE (the package re-exports the function `fit`)
from distributedpytorch_b200.trainer.steppers import make_stepper
from distributedpytorch_b200.utils import data as D

pytestmark = pytest.mark.gpu
H, W = 64, 96


def _loader(n, batch, drop_last):
    ds = D.SyntheticSegmentation(n, (W, H), 0)
    return DataLoader(ds, batch_size=batch, shuffle=False, drop_last=drop_last)


@pytest.mark.parametrize("variant", ["ref", "milesial"])
def test_loss_curve_is_the_mean_of_ten_distinct_step_losses(variant, tmp_path, monkeypatch):
    monkeypatch.chdir(tmp_path)
    torch.manual_seed(0)
    model = UNet() if variant == "ref" else UNet(variant="milesial", n_classes=2)
    stepper, native = make_stepper(model, None, 2, (W, H), "cuda:0", lr=1e-3)
    assert native
    seen = []
    inner = stepper.train_step

    def spy(images, masks):
        loss = inner(images, masks)
        seen.append(float(loss.item()))  # what the step really returned, read before the next step overwrites the buffer
        return loss

    stepper.train_step = spy
    rows, _ = F._train_epochs(stepper, _loader(24, 2, True), _loader(4, 2, True), 1, "singleGPU", 1e-3, model=model)
    assert len(seen) == 12 and len(set(round(v, 6) for v in seen[:10])) > 5, seen
    assert len(rows) == 1 and rows[0][0] == 10
    assert abs(rows[0][2] - float(np.mean(seen[:10]))) < 1e-5, (rows, seen)


@pytest.mark.parametrize("variant", ["ref", "milesial"])
def test_incomplete_last_batch_is_trained_on(variant):
    torch.manual_seed(0)
    model = UNet() if variant == "ref" else UNet(variant="milesial", n_classes=2)
    stepper, _ = make_stepper(model, None, 4, (W, H), "cuda:0", lr=1e-3)
    assert stepper.supports_partial_batch
    batches = list(_loader(6, 4, False))
    assert [b["image"].shape[0] for b in batches] == [4, 2]
    w = next(p for n, p in model.named_parameters() if n.endswith("weight") and p.dim() == 4)
    losses = []
    for b in batches + batches:
        before = w.detach().clone()
        m = b["mask"].float() if variant == "ref" else b["mask"].to(torch.uint8)
        losses.append(float(stepper.train_step(b["image"], m).item()))
        torch.cuda.synchronize()
        assert not torch.equal(before, w.detach()), "the step did not update the parameters"
    assert all(np.isfinite(losses)) and losses[2] < losses[0] and losses[3] < losses[1], losses


@pytest.mark.multigpu
def test_engine_runs_on_its_own_device_when_another_is_current():
    if torch.cuda.device_count() < 2:
        pytest.skip("needs two GPUs")
    from distributedpytorch_b200.engine import NativeUNetEngine

    torch.manual_seed(0)
    x = torch.rand(2, 3, H, W)
    t = (torch.rand(2, H, W) > 0.5).float()
    torch.manual_seed(1)
    m0 = UNet()
    torch.manual_seed(1)
    m1 = UNet()
    e0 = NativeUNetEngine(m0, 2, H, W, "cuda:0", lr=1e-3)
    e1 = NativeUNetEngine(m1, 2, H, W, "cuda:1", lr=1e-3)
    with torch.cuda.device(0):  # device 0 is current while the cuda:1 engine steps
        l1 = [float(e1.train_step(x.to("cuda:1"), t.to("cuda:1")).item()) for _ in range(3)]
    l0 = [float(e0.train_step(x.cuda(0), t.cuda(0)).item()) for _ in range(3)]
    assert np.allclose(l0, l1, rtol=1e-4, atol=1e-5), (l0, l1)


@pytest.mark.parametrize("variant", ["ref", "milesial"])
def test_prefetch_into_the_live_input_buffers_matches_explicit_batches(variant):
    """prefetch() copies the next batch from pinned host memory straight into the engine's input buffers while the step in
    flight computes (the copy waits for the program events "images consumed" / "target consumed"): same losses as handing
    every batch to train_step, i.e. no step ever saw a half-overwritten input."""
    def make():
        torch.manual_seed(3)
        model = UNet() if variant == "ref" else UNet(variant="milesial", n_classes=2)
        return make_stepper(model, None, 2, (W, H), "cuda:0", lr=1e-3)[0]

    g = torch.Generator().manual_seed(11)
    host = []
    for _ in range(4):
        x = torch.rand(2, 3, H, W, generator=g).pin_memory()
        m = (torch.rand(2, H, W, generator=g) > 0.5)
        host.append((x, (m.float() if variant == "ref" else m.to(torch.uint8)).pin_memory()))
    a, b = make(), make()
    want = [float(a.train_step(*host[k % 4]).item()) for k in range(8)]
    got = []
    b.prefetch(*host[0])
    for k in range(8):
        loss = b.train_step()
        b.prefetch(*host[(k + 1) % 4])
        got.append(float(loss.item()))
    assert np.allclose(got, want, rtol=1e-5, atol=1e-6), (got, want)
