"""Convergence of the native bf16 engines against the same module on fp32 torch ops, from the same initial weights, on the
procedural segmentation data (the north star's "Dice-equivalent convergence on synthetic data").

    python scripts/convergence.py --variant ref --size 96x64 --steps 300 --out profiles/convergence_ref_96x64.json
    python scripts/convergence.py --variant milesial --size 960x640 --steps 300 --batch 4

Both arms see the same batches in the same order.  Recorded: training loss every 10 steps, and at the end the mean
validation loss (the reference's evaluate.py metric) and the Dice score of the hard masks on held-out images.
"""
import argparse
import copy
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.trainer.steppers import MaskAdapter, TorchStepper, make_stepper  # noqa: E402
from distributedpytorch_b200.utils import data as D  # noqa: E402
from distributedpytorch_b200.utils.loss import CEDiceLoss, Loss, dice_score  # noqa: E402


def batches(ds, idx, batch, device):
    out = []
    for i in range(0, len(idx) - batch + 1, batch):
        items = [ds[j] for j in idx[i:i + batch]]
        out.append((torch.stack([it["image"] for it in items]).to(device), torch.stack([it["mask"] for it in items]).to(device)))
    return out


def run(variant="ref", size=(96, 64), steps=300, batch=4, lr=1e-3, n_train=32, n_val=8, seed=0, device="cuda:0", log=None):
    W, H = size
    torch.manual_seed(seed)
    base = UNet() if variant == "ref" else UNet(variant="milesial", n_classes=2)
    crit = Loss() if variant == "ref" else CEDiceLoss(2)
    n_classes = 1 if variant == "ref" else 2
    opt = "adam"
    ds = D.SyntheticSegmentation(n_train + n_val, (W, H), seed)
    train = batches(ds, list(range(n_train)), batch, device)
    val = batches(ds, list(range(n_train, n_train + n_val)), batch, device)
    native, is_native = make_stepper(copy.deepcopy(base), crit, batch, (W, H), device, lr=lr, optimizer=opt)
    assert is_native, "the native library must be loaded"
    oracle_model = copy.deepcopy(base).to(device)
    oracle = TorchStepper(oracle_model, MaskAdapter(crit, binary_prob=variant == "ref"), lr=lr, optimizer=opt,
                          loss_scale=float(batch), device=device)
    arms = {"native_bf16": native, "torch_fp32": oracle}
    curves = {k: [] for k in arms}
    for name, st in arms.items():
        window = []
        for s in range(steps):
            x, m = train[s % len(train)]
            if name == "native_bf16":
                m = m.float() if getattr(st, "wants_float_masks", True) else m.to(torch.uint8)
            window.append(st.train_step(x, m).detach().clone())
            if (s + 1) % 10 == 0:
                curves[name].append([s + 1, float(torch.stack(window).float().mean())])
                window = []
    res = {"variant": variant, "size": [W, H], "steps": steps, "batch": batch, "lr": lr, "optimizer": opt, "curves": curves}
    for name, st in arms.items():
        mod = getattr(st, "model", None)
        if mod is not None:
            mod.eval()
        losses, dices = [], []
        for x, m in val:
            mm = m
            if name == "native_bf16":
                mm = m.float() if getattr(st, "wants_float_masks", True) else m.to(torch.uint8)
            losses.append(float(st.evaluate_batch(x, mm)))
            dices.append(float(dice_score(st.forward(x).float().cpu(), m.cpu(), n_classes)))
        if mod is not None:
            mod.train()
        res[name] = {"val_loss": sum(losses) / len(losses), "dice": sum(dices) / len(dices),
                     "first_train_loss": curves[name][0][1], "last_train_loss": curves[name][-1][1]}
    if log:
        os.makedirs(os.path.dirname(log) or ".", exist_ok=True)
        json.dump(res, open(log, "w"), indent=1)
    return res


if __name__ == "__main__":
    ap = argparse.ArgumentParser()
    ap.add_argument("--variant", default="ref", choices=["ref", "milesial"])
    ap.add_argument("--size", default="96x64", help="WxH")
    ap.add_argument("--steps", type=int, default=300)
    ap.add_argument("--batch", type=int, default=4)
    ap.add_argument("--lr", type=float, default=1e-3)
    ap.add_argument("--out", default=None)
    a = ap.parse_args()
    w, h = (int(v) for v in a.size.lower().split("x"))
    r = run(a.variant, (w, h), a.steps, a.batch, a.lr, log=a.out)
    print(json.dumps({k: v for k, v in r.items() if k != "curves"}))
