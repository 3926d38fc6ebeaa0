"""tcgen05 implicit-GEMM convolution kernels vs plain PyTorch fp32 (same op), on a real B200."""
import pytest
import torch

from distributedpytorch_b200.ops import native as N
from distributedpytorch_b200.ops import reference as R
from gpu_util import assert_close, rand_bf16, rand_f32

pytestmark = pytest.mark.gpu


def _run_conv3x3(mode, B, H, W, cins, cout, *, relu=True, bias=True, tile_s=0, tile_n=0, cin_real=None):
    srcs = [rand_bf16(B, H, W, c, seed=11 + i) for i, c in enumerate(cins)]
    cin_total = sum(cins)
    cin_w = cin_real or cin_total
    w = rand_f32(cout, cin_w, 3, 3, scale=(9 * cin_w) ** -0.5 * 2, seed=3)
    w = w.to(torch.bfloat16).float()
    b = rand_f32(cout, seed=4) if bias else None
    wf = R.conv3x3_weight_fwd(w, cin_total)
    out = torch.full((B, H, W, cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=mode, batch=B, h=H, w=W, srcs=srcs, weight=wf, cout=cout, dsts=[out], bias=b, relu=relu,
                      tile_s=tile_s, tile_n=tile_n)
    plan.launch()
    torch.cuda.synchronize()
    ref = R.conv3x3(srcs, w, b, relu)
    assert_close(out, ref, 2e-2, 2e-2, f"conv3x3 mode={mode} {cins}->{cout} {B}x{H}x{W} info={plan.info()}")


# The aligned one-box-per-tap path first: it only needs canonical UMMA layouts.
@pytest.mark.parametrize("cins,cout", [((64,), 64), ((32,), 32), ((16,), 32), ((128,), 128), ((64,), 256)])
def test_conv3x3_tapbox(cins, cout):
    _run_conv3x3(N.CONV_3X3_TAPBOX, 2, 40, 60, cins, cout)


# The halo path: one box per K chunk, nine shifted descriptors.
@pytest.mark.parametrize("cins,cout,tile_s", [
    ((64,), 64, 1), ((64,), 64, 2), ((64,), 64, 4),
    ((32,), 32, 4), ((32,), 32, 1),
    ((16,), 32, 4),
    ((128,), 128, 2), ((256,), 256, 1),
    ((32, 32), 32, 4), ((64, 64), 64, 2), ((128, 128), 128, 1),
    ((512,), 512, 1),
])
def test_conv3x3_halo(cins, cout, tile_s):
    _run_conv3x3(N.CONV_3X3, 2, 40, 60, cins, cout, tile_s=tile_s)


def test_conv3x3_first_layer_padded_channels():
    # 3 real channels padded to 16 (what dpt_pack_input produces)
    B, H, W = 2, 48, 64
    x3 = rand_f32(B, 3, H, W, seed=5)
    x16 = torch.empty(B, H, W, 16, dtype=torch.bfloat16, device="cuda")
    N.pack_input(x3, x16)
    torch.cuda.synchronize()
    ref16 = torch.zeros(B, H, W, 16, device="cuda")
    ref16[..., :3] = x3.permute(0, 2, 3, 1)
    assert_close(x16, ref16, 1e-2, 1e-2, "pack_input")
    w = rand_f32(32, 3, 3, 3, scale=0.3, seed=6).to(torch.bfloat16).float()
    b = rand_f32(32, seed=7)
    out = torch.empty(B, H, W, 32, dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[x16], weight=R.conv3x3_weight_fwd(w, 16), cout=32,
                      dsts=[out], bias=b, relu=True)
    plan.launch()
    torch.cuda.synchronize()
    assert_close(out, R.conv3x3([x16], w, b, True), 2e-2, 2e-2, "first layer")


def test_conv3x3_full_res_many_tiles():
    # more tiles than SMs, persistent loop with several tiles per CTA, double-buffered accumulators
    _run_conv3x3(N.CONV_3X3, 2, 160, 240, (32,), 32)
    _run_conv3x3(N.CONV_3X3, 1, 160, 240, (64,), 128)


def test_conv3x3_dgrad_split_and_mask():
    # backward-data of a conv whose input was cat(skip[32], up[32]): two destinations, ReLU mask on one
    B, H, W, cout, cin = 2, 48, 64, 32, 64
    dz = rand_bf16(B, H, W, cout, seed=21)
    w = rand_f32(cout, cin, 3, 3, scale=0.1, seed=22).to(torch.bfloat16).float()
    act = rand_bf16(B, H, W, 32, seed=23)  # saved activation for the mask of destination 1
    d0 = torch.empty(B, H, W, 32, dtype=torch.bfloat16, device="cuda")
    d1 = torch.empty(B, H, W, 32, dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[dz], weight=R.conv3x3_weight_bwd(w), cout=cin,
                      dsts=[d0, d1], masks=[None, act])
    plan.launch()
    torch.cuda.synchronize()
    ref = R.conv3x3_dgrad(dz, w)
    assert_close(d0, ref[..., :32], 2e-2, 2e-2, "dgrad dst0")
    assert_close(d1, ref[..., 32:] * (act.float() > 0), 2e-2, 2e-2, "dgrad dst1 (masked)")


@pytest.mark.parametrize("cin,cout", [(64, 32), (128, 64), (512, 256)])
def test_convT_forward_pixel_shuffle(cin, cout):
    B, H, W = 2, 24, 40
    x = rand_bf16(B, H, W, cin, seed=31)
    w = rand_f32(cin, cout, 2, 2, scale=cin ** -0.5 * 2, seed=32).to(torch.bfloat16).float()
    b = rand_f32(cout, seed=33)
    out = torch.full((B, 2 * H, 2 * W, cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_1X1, out_mode=N.OUT_SHUFFLE2, batch=B, h=H, w=W, srcs=[x],
                      weight=R.convT_weight_fwd(w), cout=4 * cout, dsts=[out], bias=b.repeat(4).contiguous())
    plan.launch()
    torch.cuda.synchronize()
    assert_close(out, R.convT2x2(x, w, b), 2e-2, 2e-2, f"convT fwd {cin}->{cout} info={plan.info()}")


@pytest.mark.parametrize("cin,cout", [(64, 32), (256, 128)])
def test_convT_dgrad_strided_gather(cin, cout):
    B, H, W = 2, 24, 40
    dy = rand_bf16(B, 2 * H, 2 * W, cout, seed=41)
    w = rand_f32(cin, cout, 2, 2, scale=cout ** -0.5, seed=42).to(torch.bfloat16).float()
    act = rand_bf16(B, H, W, cin, seed=43)
    dx = torch.empty(B, H, W, cin, dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_2X2S2, batch=B, h=H, w=W, srcs=[dy], weight=R.convT_weight_bwd(w), cout=cin,
                      dsts=[dx], masks=[act])
    plan.launch()
    torch.cuda.synchronize()
    ref = R.convT2x2_dgrad(dy, w) * (act.float() > 0)
    assert_close(dx, ref, 2e-2, 2e-2, f"convT dgrad {cin}<-{cout} info={plan.info()}")


@pytest.mark.parametrize("B,H,W,cin,cout", [(2, 40, 60, 64, 64), (1, 48, 64, 32, 32), (2, 32, 48, 64, 128),
                                             (1, 40, 60, 128, 256), (1, 16, 24, 256, 1024), (2, 64, 96, 16, 64)])
def test_conv_bn_stats(B, H, W, cin, cout):
    """BatchNorm statistics from the conv epilogue: per-CTA rows, summed by dpt_bn_finalize -> mean / invstd; tiles that
    overhang the image (40x60), one / two / eight N tiles, NT = 32 / 64 / 128; and bit-reproducibility."""
    x = rand_bf16(B, H, W, cin, seed=51)
    w = rand_f32(cout, cin, 3, 3, scale=(9 * cin) ** -0.5 * 2, seed=52).to(torch.bfloat16).float()
    rows = torch.zeros(148, 2, cout, device="cuda")
    out = torch.empty(B, H, W, cout, dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[x], weight=R.conv3x3_weight_fwd(w), cout=cout, dsts=[out],
                      stats=rows)
    info = plan.info()
    plan.launch()
    torch.cuda.synchronize()
    ref = R.conv3x3([x], w, None, False)
    assert_close(out, ref, 2e-2, 2e-2, f"conv output {info}")
    tot = rows[:info["grid"]].sum(0)  # rows a CTA does not own stay zero in this zero-initialised buffer
    s1, s2 = ref.sum(dim=(0, 1, 2)), (ref * ref).sum(dim=(0, 1, 2))
    assert_close(tot[0], s1, 1e-3 * float(s2.max().sqrt()) * (B * H * W) ** 0.5 + 0.05, 1e-3, f"bn sum {info}")
    assert_close(tot[1], s2, 1e-3, 2e-3, f"bn sum of squares {info}")
    # finalize from the rows (garbage in the rows a channel's CTAs do not own must not matter)
    rows2 = torch.full((148, 2, cout), float("nan"), device="cuda")
    plan2 = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[x], weight=R.conv3x3_weight_fwd(w), cout=cout, dsts=[out],
                       stats=rows2)
    plan2.launch()
    gamma, beta = torch.ones(cout, device="cuda"), torch.zeros(cout, device="cuda")
    rm, rv = torch.zeros(cout, device="cuda"), torch.ones(cout, device="cuda")
    bn = torch.empty(4, cout, device="cuda")
    nt = info["tile_n"]
    P = B * H * W
    N.check(N.lib().dpt_bn_finalize(rows2.data_ptr(), info["grid"] // (cout // nt), nt, gamma.data_ptr(), beta.data_ptr(),
                                    rm.data_ptr(), rv.data_ptr(), bn.data_ptr(), P, cout, 1e-5, 0.1, N.stream_ptr()))
    torch.cuda.synchronize()
    mean, var = s1 / P, s2 / P - (s1 / P) ** 2
    assert_close(bn[0], mean, 2e-3, 2e-3, "mean")
    assert_close(bn[1], (var + 1e-5).rsqrt(), 1e-3, 3e-3, "invstd")
    first = rows2.clone()
    plan2.launch()
    torch.cuda.synchronize()
    assert torch.equal(torch.nan_to_num(first), torch.nan_to_num(rows2)), "statistics rows are not bit-reproducible"


@pytest.mark.parametrize("cluster", [2, 4])
@pytest.mark.parametrize("B,H,W,cins,cout,stats", [(1, 40, 56, (128,), 128, False), (3, 16, 24, (64, 64), 256, False),
                                                    (1, 48, 40, (256,), 512, True), (2, 16, 40, (128,), 128, True)])
def test_cluster_multicast_filter_tiles_match_the_plain_kernel(cluster, B, H, W, cins, cout, stats):
    """Thread-block clusters that share every streamed filter tile through TMA multicast (wide 3x3 layers): same result,
    bit for bit, as the one-CTA kernel - odd numbers of pixel tiles (a short last group), concat sources, several N tiles,
    and the BatchNorm-statistics rows."""
    srcs = [rand_bf16(B, H, W, c, seed=61 + i) for i, c in enumerate(cins)]
    cin = sum(cins)
    w = rand_f32(cout, cin, 3, 3, scale=(9 * cin) ** -0.5 * 2, seed=62).to(torch.bfloat16).float()
    wf = R.conv3x3_weight_fwd(w)
    outs, rows = [], []
    for cl in (1, cluster):
        out = torch.zeros(B, H, W, cout, dtype=torch.bfloat16, device="cuda")
        st = torch.zeros(148, 2, cout, device="cuda") if stats else None
        plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=srcs, weight=wf, cout=cout, dsts=[out], stats=st, cluster=cl)
        info = plan.info()
        if cl > 1 and info["cluster"] != cl:
            pytest.skip(f"no cluster-{cl} kernel for this shape: {info}")
        plan.launch()
        plan.launch()
        torch.cuda.synchronize()
        outs.append(out)
        rows.append(st[:info["grid"]].sum(0) if stats else None)
    assert_close(outs[1], R.conv3x3(srcs, w, None, False), 2e-2, 2e-2, f"cluster {cluster} conv {info}")
    assert torch.equal(outs[0], outs[1]), "cluster kernel differs from the one-CTA kernel"
    if stats:
        assert_close(rows[1], rows[0], 1e-2, 1e-4, "BatchNorm statistics rows")


# The encoder's MaxPool2d(2, 2) written by the conv epilogue itself: the pooled tensor must be exactly the maximum of
# the stored bf16 activation (same rounding), also with tiles that hang over the image edge.
@pytest.mark.parametrize("B,H,W,cin,cout", [(2, 40, 60, 32, 32), (1, 48, 64, 64, 64), (3, 20, 36, 64, 128), (2, 16, 24, 128, 256),
                                            (1, 10, 14, 256, 512)])
def test_conv3x3_fused_maxpool_output(B, H, W, cin, cout):
    x = rand_bf16(B, H, W, cin, seed=21)
    w = rand_f32(cout, cin, 3, 3, scale=(9 * cin) ** -0.5 * 2, seed=22).to(torch.bfloat16).float()
    b = rand_f32(cout, seed=23)
    wf = R.conv3x3_weight_fwd(w, cin)
    out = torch.full((B, H, W, cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    pooled = torch.full((B, H // 2, W // 2, cout), float("nan"), dtype=torch.bfloat16, device="cuda")
    plan = N.ConvPlan(mode=N.CONV_3X3, batch=B, h=H, w=W, srcs=[x], weight=wf, cout=cout, dsts=[out], bias=b, relu=True,
                      pooled=pooled)
    plan.launch()
    torch.cuda.synchronize()
    assert_close(out, R.conv3x3([x], w, b, True), 2e-2, 2e-2, f"conv with pooled output {plan.info()}")
    want = torch.nn.functional.max_pool2d(out.float().permute(0, 3, 1, 2), 2, 2).permute(0, 2, 3, 1)
    assert torch.equal(pooled.float(), want), f"fused max-pool differs from max_pool2d of the stored activation {plan.info()}"


def test_conv3x3_fused_maxpool_rejects_odd_extent():
    x = rand_bf16(1, 17, 24, 32, seed=1)
    wf = R.conv3x3_weight_fwd(rand_f32(32, 32, 3, 3, seed=2), 32)
    out = torch.empty(1, 17, 24, 32, dtype=torch.bfloat16, device="cuda")
    pooled = torch.empty(1, 8, 12, 32, dtype=torch.bfloat16, device="cuda")
    with pytest.raises(RuntimeError, match="even h and w"):
        N.ConvPlan(mode=N.CONV_3X3, batch=1, h=17, w=24, srcs=[x], weight=wf, cout=32, dsts=[out], bias=rand_f32(32, seed=3),
                   relu=True, pooled=pooled)
