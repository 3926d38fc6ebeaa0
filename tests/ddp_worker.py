"""2-rank worker for the native DDP test (launched with torch.distributed.run by test_gpu_ddp.py)."""
import copy
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.parallel.ddp import FusedDDP  # noqa: E402
from distributedpytorch_b200.utils.loss import Loss  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, H, W = 2, 64, 96
    torch.manual_seed(100 + rank)  # different init per rank: the constructor must broadcast rank 0's
    model = UNet()
    if os.environ.get("DPT_ALGO") == "multimem":
        from distributedpytorch_b200.parallel.comm import multicast_supported

        if not multicast_supported(torch.device("cuda", local)):  # no NVSwitch multicast on this box: nothing to test
            print("DDPSKIP multicast unsupported", flush=True)
            dist.destroy_process_group()
            return
    ddp = FusedDDP(model, B, H, W, f"cuda:{local}", lr=1e-3, algo=os.environ.get("DPT_ALGO", "two_shot"),
                   overlap=bool(int(os.environ.get("DPT_DDP_OVERLAP", "0"))))
    # oracle: torch DDP semantics by hand on fp32 torch ops, starting from the broadcast parameters
    ref = copy.deepcopy(model).float()
    for p in ref.parameters():
        p.data = p.data.clone()
    w0 = [p.detach().clone() for p in ref.parameters()]
    gathered = [torch.empty_like(w0[0]) for _ in range(world)]
    dist.all_gather(gathered, w0[0])
    same_init = all(torch.equal(gathered[0], g) for g in gathered)
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3, weight_decay=1e-8)
    g = torch.Generator(device="cuda").manual_seed(7 + rank)  # different data per rank
    losses, losses_ref = [], []
    for step in range(4):
        x = torch.rand(B, 3, H, W, device="cuda", generator=g)
        t = (torch.rand(B, H, W, device="cuda", generator=g) > 0.5).float()
        loss_ref = Loss()(ref(x), t.unsqueeze(1))
        opt.zero_grad()
        (B * loss_ref).backward()
        for p in ref.parameters():
            dist.all_reduce(p.grad)
            p.grad /= world
        opt.step()
        losses_ref.append(loss_ref.item())
        losses.append(ddp.train_step(x, t).item())
    torch.cuda.synchronize()
    # replicas must hold identical parameters after the fused all-gather
    flat = ddp.engine.arena.params.clone()
    other = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(other, flat)
    identical = all(torch.equal(other[0], o) for o in other)
    cos = []
    for p, q in zip(model.parameters(), ref.parameters()):
        a, b = (p.detach() - 0).flatten(), q.detach().flatten()
        cos.append(float(torch.dot(a, b) / (a.norm() * b.norm() + 1e-30)))
    upd = []
    for p, q, w in zip(model.parameters(), ref.parameters(), w0):
        da, db = (p.detach() - w).flatten(), (q.detach() - w).flatten()
        upd.append(float(torch.dot(da, db) / (da.norm() * db.norm() + 1e-30)))
    out = {"rank": rank, "same_init": same_init, "identical_replicas": identical, "min_param_cos": min(cos),
           "mean_update_cos": sum(upd) / len(upd), "losses": losses, "losses_ref": losses_ref}
    print("DDPRESULT " + json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
