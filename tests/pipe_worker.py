"""N-stage pipeline worker on CPU/gloo (or CUDA/nccl): pipeline result must equal the un-pipelined model."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.parallel import pipe as P  # noqa: E402
from distributedpytorch_b200.utils.loss import CEDiceLoss, Loss  # noqa: E402


def main():
    rank, world = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"])
    schedule = os.environ.get("DPT_SCHEDULE", "gpipe")
    n_micro = int(os.environ.get("DPT_MICRO", "2"))
    backend = os.environ.get("DPT_BACKEND", "gloo")
    if backend == "nccl":
        torch.cuda.set_device(int(os.environ["LOCAL_RANK"]))
        dist.init_process_group(backend)  # lazy init: per-pair P2P communicators (see make_pair_groups)
    else:
        dist.init_process_group(backend)
    milesial = os.environ.get("DPT_VARIANT", "ref") == "milesial"
    B, H, W = 4, 32, 48
    torch.manual_seed(0)
    model = UNet(variant="milesial", n_classes=2) if milesial else UNet()
    ref = copy.deepcopy(model)
    stage_of = P.partition([c for _, c in model.stage_costs()], world)
    dev = torch.device("cuda", int(os.environ["LOCAL_RANK"])) if backend == "nccl" else torch.device("cpu")
    groups = P.make_pair_groups(stage_of, list(range(world)), device=dev)
    lo = stage_of.index(rank)
    hi = len(stage_of) - stage_of[::-1].index(rank)
    sizes = P.micro_batch_sizes(B, n_micro)
    if backend == "nccl":
        from distributedpytorch_b200.engine.stage_engine import NativeStage

        local = int(os.environ["LOCAL_RANK"])
        torch.cuda.set_device(local)
        be = NativeStage(model, lo, hi, sizes, H, W, f"cuda:{local}", lr=1e-3, loss_scale=float(B),
                         whole_batch=schedule == "gpipe", optimizer="adam")
    else:
        be = P.TorchStage(model, lo, hi, sizes, H, W, "cpu", criterion=Loss(), lr=1e-3, loss_scale=float(B),
                          whole_batch=schedule == "gpipe")
    stage = P.PipelineStage(be, rank, stage_of, len(sizes), schedule, groups)
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3, weight_decay=1e-8)
    g = torch.Generator().manual_seed(3)
    losses, losses_ref = [], []
    yy, xx = torch.meshgrid(torch.arange(H), torch.arange(W), indexing="ij")
    blob = (((yy - H / 2) ** 2 / (H / 3) ** 2 + (xx - W / 2) ** 2 / (W / 3) ** 2) < 1).long().expand(B, H, W).contiguous()
    for step in range(3):
        x = torch.rand(B, 3, H, W, generator=g)
        t = (torch.rand(B, H, W, generator=g) > 0.5).long()
        if milesial:
            t = blob  # a learnable target: the loss must also go down
        off = 0
        for mb, b in enumerate(sizes):
            tm = t[off:off + b].float() if backend == "nccl" else t[off:off + b]
            if milesial:
                tm = t[off:off + b].to(torch.uint8)
            be.load_micro_batch(mb, x[off:off + b] if rank == 0 else None, tm if rank == world - 1 else None)
            off += b
        loss = stage.train_step()
        if milesial:
            # oracle: every micro-batch goes through the module on its own (BatchNorm statistics per micro-batch), the
            # loss is taken over the concatenated logits of the whole batch - under GPipe and 1F1B alike
            ref.train()
            outs, off = [], 0
            for b in sizes:
                outs.append(ref(x[off:off + b]))
                off += b
            lr_ = CEDiceLoss(2)(torch.cat(outs, 0), t)
        elif schedule == "gpipe":
            lr_ = Loss()(ref(x), t.float().unsqueeze(1))
        else:  # per-micro-batch loss statistics
            lr_, off = 0, 0
            for b in sizes:
                lr_ = lr_ + Loss()(ref(x[off:off + b]), t[off:off + b].float().unsqueeze(1)) * (b / B)
                off += b
        opt.zero_grad()
        (B * lr_).backward()
        opt.step()
        losses_ref.append(float(lr_))
        if loss is not None:
            losses.append(float(loss))
    # every stage checks the parameters it owns against the un-pipelined run
    worst, mean_num, mean_den = 0.0, 0.0, 0
    ref_sd = dict(ref.named_parameters())
    names = be.owned_parameter_names() if backend == "nccl" else None
    owned = (lambda n: n in names) if backend == "nccl" else be._owns
    for n, p in model.named_parameters():
        if owned(n):
            d = (p.detach().cpu() - ref_sd[n].detach()).abs()
            worst = max(worst, float(d.max()))
            mean_num += float(d.sum())
            mean_den += d.numel()
    if backend == "nccl":
        torch.cuda.synchronize()
    print("PIPERESULT " + json.dumps({"rank": rank, "stage_of": stage_of, "losses": losses, "losses_ref": losses_ref,
                                      "max_param_diff": worst, "mean_param_diff": mean_num / max(mean_den, 1)}), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
