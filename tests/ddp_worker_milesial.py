"""2-rank worker: FusedDDP on the BatchNorm UNet - replicas stay identical, BN buffers follow rank 0, loss goes down."""
import json
import os
import sys

import torch
import torch.distributed as dist

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.models import UNet  # noqa: E402
from distributedpytorch_b200.parallel.ddp import FusedDDP  # noqa: E402


def main():
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    B, H, W = 2, 64, 96
    torch.manual_seed(100 + rank)
    model = UNet(variant="milesial", n_classes=2)
    algo = os.environ.get("DPT_ALGO", "auto")
    if algo == "multimem":
        from distributedpytorch_b200.parallel.comm import multicast_supported

        if not multicast_supported(torch.device("cuda", local)):
            print("DDPSKIP multicast unsupported", flush=True)
            dist.destroy_process_group()
            return
    ddp = FusedDDP(model, B, H, W, f"cuda:{local}", lr=1e-3, optimizer="adam", algo=algo)
    g = torch.Generator(device="cuda").manual_seed(7 + rank)
    yy, xx = torch.meshgrid(torch.arange(H, device="cuda"), torch.arange(W, device="cuda"), indexing="ij")
    t = (((yy - H / 2) ** 2 / (H / 3) ** 2 + (xx - W / 2) ** 2 / (W / 3) ** 2) < 1).to(torch.uint8).expand(B, H, W).contiguous()
    losses = []
    for _ in range(8):
        x = torch.rand(B, 3, H, W, device="cuda", generator=g)
        losses.append(ddp.train_step(x, t).item())
    torch.cuda.synchronize()
    flat = ddp.engine.params.clone()
    other = [torch.empty_like(flat) for _ in range(world)]
    dist.all_gather(other, flat)
    rm = dict(model.named_buffers())["inc.double_conv.1.running_mean"].clone()
    rms = [torch.empty_like(rm) for _ in range(world)]
    dist.all_gather(rms, rm)
    out = {"rank": rank, "identical_replicas": all(torch.equal(other[0], o) for o in other),
           "identical_bn_buffers": all(torch.equal(rms[0], r) for r in rms), "bn_moved": float(rm.abs().sum()) > 0, "losses": losses}
    print("DDPRESULT " + json.dumps(out), flush=True)
    dist.barrier()
    dist.destroy_process_group()


if __name__ == "__main__":
    main()
