"""Dice-equivalent convergence: the native bf16 engines and the same module on fp32 torch ops, trained from the same
initial weights on the same synthetic batches, end at the same validation loss (5 %) and Dice score (0.02)."""
import os
import sys

import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "scripts"))


def _check(res):
    n, t = res["native_bf16"], res["torch_fp32"]
    assert n["last_train_loss"] < 0.5 * n["first_train_loss"], res
    assert abs(n["val_loss"] - t["val_loss"]) < 0.05 * abs(t["val_loss"]) + 0.02, (n, t)
    assert abs(n["dice"] - t["dice"]) < 0.02 and n["dice"] > 0.9, (n, t)


@pytest.mark.parametrize("variant", ["ref", "milesial"])
def test_convergence_matches_fp32_torch_small(variant):
    import convergence

    _check(convergence.run(variant, (96, 64), steps=300, batch=4, lr=1e-3))


def test_convergence_matches_fp32_torch_at_the_benchmark_shape():
    import convergence

    out = os.environ.get("DPT_CONVERGENCE_OUT")
    _check(convergence.run("ref", (960, 640), steps=300, batch=4, lr=1e-3, n_train=16, n_val=4, log=out))
