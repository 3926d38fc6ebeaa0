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


def _algo() -> str:
    for a in sys.argv[1:]:
        if a.startswith("--algo="):
            return a.split("=", 1)[1]
    return "auto"


def main_milesial():
    """BatchNorm variant: replicas stay identical, the first loss equals the fp32 oracle with PER-REPLICA batch statistics
    and a whole-batch loss (torch DataParallel semantics), the loss goes down, evaluation uses the root's running stats."""
    from distributedpytorch_b200.parallel.dp import chunk_sizes
    from distributedpytorch_b200.utils.loss import CEDiceLoss

    n = torch.cuda.device_count()
    B, H, W = 4, 64, 96
    torch.manual_seed(0)
    base = UNet(variant="milesial", n_classes=2)
    oracle = copy.deepcopy(base).cuda(0).train()
    dp = NativeDP(copy.deepcopy(base), B, H, W, list(range(n)), lr=1e-3, optimizer="adam", algo=_algo())
    g = torch.Generator(device="cuda:0").manual_seed(5)
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda:0"), torch.arange(W, device="cuda:0"), indexing="ij")
    t = (((yy - H / 2) ** 2 / (H / 3) ** 2 + (xx - W / 2) ** 2 / (W / 3) ** 2) < 1).long().expand(B, H, W).contiguous()
    x = torch.rand(B, 3, H, W, device="cuda:0", generator=g)
    outs, off = [], 0
    with torch.no_grad():
        for b in chunk_sizes(B, n):
            outs.append(oracle(x[off:off + b]))
            off += b
        loss_ref = float(CEDiceLoss(2)(torch.cat(outs, 0), t))
    losses = [dp.train_step(x, t.to(torch.uint8)).item() for _ in range(8)]
    dp.synchronize()
    identical = all(torch.equal(dp.engines[0].params.cpu(), e.params.cpu()) for e in dp.engines)
    logits = dp.forward(x).clone()
    dp.synchronize()
    root = dp.engines[0].model
    root.eval()
    with torch.no_grad():
        lg_ref = root(x)  # the root module's parameters and buffers ARE what the engines trained
    rel = float((logits - lg_ref).abs().max() / (lg_ref.abs().max() + 1e-6))
    val = float(dp.evaluate_batch(x, t))
    out = {"gpus": n, "algo": dp.algo, "losses": losses, "loss_ref_step0": loss_ref, "replicas_identical": identical, "eval_rel_diff": rel,
           "eval_loss": val}
    print("DPRESULT " + json.dumps(out))
    ok = (identical and abs(losses[0] - loss_ref) < 0.03 * abs(loss_ref) + 5e-3 and losses[-1] < losses[0] and rel < 0.15
          and val == val)
    sys.exit(0 if ok else 1)


def main():
    if len(sys.argv) > 1 and sys.argv[1] == "milesial":
        return main_milesial()
    n = torch.cuda.device_count()
    B, H, W = 4, 64, 96
    torch.manual_seed(0)
    base = UNet()
    single_model = copy.deepcopy(base)
    single = NativeUNetEngine(single_model, B, H, W, "cuda:0", lr=1e-3)
    dp_model = copy.deepcopy(base)
    dp = NativeDP(dp_model, B, H, W, list(range(n)), lr=1e-3, algo=_algo())
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
    out = {"gpus": n, "algo": dp.algo, "losses_single": l1, "losses_dp": l2, "max_param_diff_vs_single": max(diffs),
           "replicas_identical": identical, "max_prob_diff": float((prob - prob_single).abs().max())}
    print("DPRESULT " + json.dumps(out))
    ok = identical and max(abs(a - b) for a, b in zip(l1, l2)) < 2e-3 and max(diffs) < 5e-3 and out["max_prob_diff"] < 2e-2
    sys.exit(0 if ok else 1)


if __name__ == "__main__":
    main()
