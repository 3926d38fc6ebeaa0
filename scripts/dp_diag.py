"""DataParallel sanity run on every visible GPU: per-step loss, wall-clock ms and whether the replicas stay bit-identical.

    python scripts/dp_diag.py two_shot|one_shot|multimem|auto HEIGHT WIDTH
"""
import sys, os, time, torch
sys.path.insert(0, os.getcwd())
from distributedpytorch_b200.models import UNet
from distributedpytorch_b200.parallel.dp import NativeDP
algo = sys.argv[1]; H, W = int(sys.argv[2]), int(sys.argv[3])
n = torch.cuda.device_count(); B = 4 * n
torch.manual_seed(0)
dp = NativeDP(UNet(), B, H, W, list(range(n)), lr=1e-4, algo=algo)
x = torch.rand(B, 3, H, W, device="cuda:0"); t = (torch.rand(B, H, W, device="cuda:0") > 0.5).float()
dp.load_batch(x, t)
for i in range(6):
    t0 = time.time(); l = dp.train_step(); dp.synchronize(); dt = time.time() - t0
    for d in range(n): torch.cuda.synchronize(d)
    print(algo, H, W, "step", i, "loss", float(l), "ms", round(dt * 1e3, 3), "ident", all(torch.equal(dp.engines[0].arena.params.cpu(), e.arena.params.cpu()) for e in dp.engines), flush=True)
