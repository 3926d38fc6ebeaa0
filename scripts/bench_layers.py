"""Per-layer device timings of the tcgen05 conv kernels at the benchmark shape (B x 640 x 960), with the
roofline each layer is judged against: max(FLOPs / measured bf16 peak, algorithmic bytes / measured HBM BW).
Writes JSON lines to stdout.  Usage: python scripts/bench_layers.py [--batch 4] [--iters 10]"""
import argparse
import json
import os
import sys

import torch

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from distributedpytorch_b200.ops import native as N  # noqa: E402
from distributedpytorch_b200.ops import reference as R  # noqa: E402

ap = argparse.ArgumentParser()
ap.add_argument("--batch", type=int, default=4)
ap.add_argument("--iters", type=int, default=10)
ap.add_argument("--only", type=str, default="")
ap.add_argument("--tile-s", type=int, default=0)
ap.add_argument("--tile-n", type=int, default=0)
ap.add_argument("--kinds", type=str, default="fprop,wgrad,dgrad")
args = ap.parse_args()

peaks = {"hbm_gbs": 6540.8, "bf16_tflops": 1650.6}
try:
    peaks.update(json.load(open(os.path.join(os.path.dirname(__file__), "..", "MEASURED_PEAKS.json"))))
except Exception:
    pass

B = args.batch
H0, W0 = 640, 960
# (name, level, cins, cout)
LAYERS = [
    ("enc1.0", 0, (16,), 32), ("enc1.2", 0, (32,), 32), ("enc2.0", 1, (32,), 64), ("enc2.2", 1, (64,), 64),
    ("enc3.0", 2, (64,), 128), ("enc3.2", 2, (128,), 128), ("enc4.0", 3, (128,), 256), ("enc4.2", 3, (256,), 256),
    ("mid.0", 4, (256,), 512), ("mid.2", 4, (512,), 512),
    ("dec1.0", 3, (256, 256), 256), ("dec1.2", 3, (256,), 256), ("dec2.0", 2, (128, 128), 128),
    ("dec2.2", 2, (128,), 128), ("dec3.0", 1, (64, 64), 64), ("dec3.2", 1, (64,), 64),
    ("dec4.0", 0, (32, 32), 32), ("dec4.2", 0, (32,), 32),
]
flush = torch.empty(256 * 1024 * 1024, dtype=torch.uint8, device="cuda")  # > 126 MB L2


def time_launch(fn, iters):
    fn()
    fn()
    fn()
    torch.cuda.synchronize()
    ts = []
    for _ in range(iters):
        flush.zero_()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        fn()
        e1.record()
        torch.cuda.synchronize()
        ts.append(e0.elapsed_time(e1))
    ts.sort()
    return ts[len(ts) // 2], ts[0]


for name, lvl, cins, cout in LAYERS:
    if args.only and not any(tok in name for tok in args.only.split(",")):
        continue
    H, W = H0 >> lvl, W0 >> lvl
    cin = sum(cins)
    srcs = [torch.randn(B, H, W, c, device="cuda").to(torch.bfloat16) for c in cins]
    wf = (torch.randn(9, cout, cin, device="cuda") * 0.05).to(torch.bfloat16)
    bias = torch.randn(cout, device="cuda")
    out = torch.empty(B, H, W, cout, dtype=torch.bfloat16, device="cuda")
    try:
        plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=srcs, weight=wf, cout=cout, dsts=[out], bias=bias,
                          relu=True, tile_s=args.tile_s, tile_n=args.tile_n)
    except N.NativeError:  # a forced tile shape that this layer cannot take (tile sweeps)
        continue
    med, best = time_launch(plan.launch, args.iters)
    px = B * H * W
    flops = 2.0 * px * cout * 9 * cin
    byts = px * (cin + cout) * 2 + wf.numel() * 2
    t_roof = max(flops / (peaks["bf16_tflops"] * 1e12), byts / (peaks["hbm_gbs"] * 1e9)) * 1e3
    rec = {"layer": name, "kind": "fprop", "shape": [B, H, W, cin, cout], "ms": round(med, 4), "ms_best": round(best, 4),
           "tflops": round(flops / med / 1e9, 1), "gbs": round(byts / med / 1e6, 1), "roof_ms": round(t_roof, 4),
           "frac_of_roof_measured": round(t_roof / med, 3), **plan.info()}
    print(json.dumps(rec), flush=True)

    # wgrad
    dz = torch.randn(B, H, W, cout, device="cuda").to(torch.bfloat16)
    dw = torch.empty(cout, cin, 3, 3, device="cuda")
    wp = N.WgradPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, z=dz, xs=srcs, dw=dw, strides=(cin * 9, 9, 1))
    med, best = time_launch(wp.launch, args.iters)
    byts = px * (cin + cout) * 2 + dw.numel() * 4
    t_roof = max(flops / (peaks["bf16_tflops"] * 1e12), byts / (peaks["hbm_gbs"] * 1e9)) * 1e3
    rec = {"layer": name, "kind": "wgrad", "shape": [B, H, W, cin, cout], "ms": round(med, 4), "ms_best": round(best, 4),
           "tflops": round(flops / med / 1e9, 1), "gbs": round(byts / med / 1e6, 1), "roof_ms": round(t_roof, 4),
           "frac_of_roof_measured": round(t_roof / med, 3), "scratch_mb": round(wp.scratch_bytes / 2**20, 1)}
    print(json.dumps(rec), flush=True)

    # dgrad: dz (cout channels) -> gradient of every source; single-source layers apply the ReLU-backward mask of
    # their input in the epilogue (what the engine does), concat layers write the two halves unmasked
    if "dgrad" in args.kinds and name != "enc1.0":
        wd = (torch.randn(9, cin, cout, device="cuda") * 0.05).to(torch.bfloat16)
        dsts = [torch.empty_like(x) for x in srcs]
        masks = [srcs[0]] if len(srcs) == 1 else [None, None]
        try:
            dp = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[dz], weight=wd, cout=cin, dsts=dsts, masks=masks,
                            tile_s=args.tile_s, tile_n=args.tile_n if cin % max(args.tile_n, 1) == 0 else 0)
        except N.NativeError:
            continue
        med, best = time_launch(dp.launch, args.iters)
        byts = px * (cin + cout + (cin if len(srcs) == 1 else 0)) * 2 + wd.numel() * 2
        t_roof = max(flops / (peaks["bf16_tflops"] * 1e12), byts / (peaks["hbm_gbs"] * 1e9)) * 1e3
        rec = {"layer": name, "kind": "dgrad", "shape": [B, H, W, cout, cin], "ms": round(med, 4), "ms_best": round(best, 4),
               "tflops": round(flops / med / 1e9, 1), "gbs": round(byts / med / 1e6, 1), "roof_ms": round(t_roof, 4),
               "frac_of_roof_measured": round(t_roof / med, 3), "masked": len(srcs) == 1, **dp.info()}
        print(json.dumps(rec), flush=True)
        del dp, wd, dsts
    del plan, wp, srcs, out, dz, dw
