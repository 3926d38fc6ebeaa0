"""Native single-process DataParallel on all visible GPUs vs the same global batch on one GPU (run on the box)."""
import copy
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.engine import NativeUNetEngine  # noqa: E402
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.parallel.dp import NativeDP  # noqa: E402


def main():
    n = torch.cuda.device_count()
    B, H, W = 4, 64, 96
    torch.manual_seed(0)
    base = UNet()
    single_model = copy.deepcopy(base)
    single = NativeUNetEngine(single_model, B, H, W, "cuda:0", lr=1e-3)
    dp_model = copy.deepcopy(base)
    dp = NativeDP(dp_model, B, H, W, list(range(n)), lr=1e-3)
    g = torch.Generator(device="cuda:0").manual_seed(5)
    l1, l2 = [], []
    for _ in range(4):
        x = torch.rand(B, 3, H, W, device="cuda:0", generator=g)
        t = (torch.rand(B, H, W, device="cuda:0", generator=g) > 0.5).float()
        l1.append(single.train_step(x, t).item())
        l2.append(dp.train_step(x, t).item())
    dp.synchronize()
    p_single = single.arena.params
    diffs = []
    for e in dp.engines:
        diffs.append(float((e.arena.params.to("cuda:0") - p_single).abs().max()))
    identical = all(torch.equal(dp.engines[0].arena.params.cpu(), e.arena.params.cpu()) for e in dp.engines)
    prob = dp.forward(x)
    dp.synchronize()
    prob_single = single.forward(x)
    torch.cuda.synchronize()
    out = {"gpus": n, "losses_single": l1, "losses_dp": l2, "max_param_diff_vs_single": max(diffs),
           "replicas_identical": identical, "max_prob_diff": float((prob - prob_single).abs().max())}
    print("DPRESULT " + json.dumps(out))
    ok = identical and max(abs(a - b) for a, b in zip(l1, l2)) < 2e-3 and max(diffs) < 5e-3 and out["max_prob_diff"] < 2e-2
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
