"""Bandwidth-bound kernels, head/loss and optimizers vs PyTorch fp32 (same op) on a real B200."""
import ctypes as C

import pytest
import torch
import torch.nn.functional as F

from distributedpytorch_b200.ops import native as N
from distributedpytorch_b200.ops import reference as R
from gpu_util import assert_close, rand_bf16, rand_f32

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("c", [32, 64, 256])
def test_maxpool(c):
    x = rand_bf16(2, 48, 64, c, seed=1)
    y = torch.empty(2, 24, 32, c, dtype=torch.bfloat16, device="cuda")
    N.maxpool2x2(x, y)
    torch.cuda.synchronize()
    assert torch.equal(y.float(), R.maxpool2x2(x))


@pytest.mark.parametrize("with_skip", [True, False])
def test_pool_bwd_fused(with_skip):
    c = 32
    act = torch.relu(rand_bf16(2, 48, 64, c, seed=2).float()).to(torch.bfloat16)
    d_pool = rand_bf16(2, 24, 32, c, seed=3)
    d_skip = rand_bf16(2, 48, 64, c, seed=4) if with_skip else None
    dz = torch.empty_like(act)
    N.pool_bwd_fused(act, d_pool, d_skip, dz, relu_mask=True)
    torch.cuda.synchronize()
    ref = R.pool_bwd_fused(act, d_pool, d_skip, True)
    assert_close(dz, ref, 1e-2, 1e-2, "pool bwd fused")
    if with_skip:  # in place on d_skip
        N.pool_bwd_fused(act, d_pool, d_skip, d_skip, relu_mask=True)
        torch.cuda.synchronize()
        assert torch.equal(d_skip, dz)


@pytest.mark.parametrize("c", [32, 64, 512])
def test_channel_sum(c):
    x = rand_bf16(3, 40, 60, c, seed=5)
    out = torch.empty(c, device="cuda")
    N.channel_sum(x, out, scale=0.5)
    torch.cuda.synchronize()
    ref = 0.5 * x.float().sum(dim=(0, 1, 2))
    assert_close(out, ref, 5e-2, 1e-3, "channel_sum")


def test_pack_input_uint8_equals_float_path():
    """Raw 8-bit pixels normalised by the packing kernel == the float tensor the reference's loader would have produced."""
    g = torch.Generator().manual_seed(5)
    u8 = torch.randint(0, 256, (2, 3, 40, 56), generator=g, dtype=torch.uint8).cuda()
    f32 = u8.float() / 255.0
    a = torch.empty(2, 40, 56, 16, dtype=torch.bfloat16, device="cuda")
    b = torch.empty_like(a)
    N.pack_input(u8, a)
    N.pack_input(f32, b)
    torch.cuda.synchronize()
    assert torch.equal(a, b)
    assert torch.equal(a[..., 3:], torch.zeros_like(a[..., 3:]))


def test_prep_weights_matches_reference_layouts():
    w3 = rand_f32(64, 32, 3, 3, seed=6)
    wt = rand_f32(64, 32, 2, 2, seed=7)
    bt = rand_f32(32, seed=8)
    w1 = rand_f32(32, 3, 3, 3, seed=9)
    wf3 = torch.empty(9, 64, 32, dtype=torch.bfloat16, device="cuda")
    wd3 = torch.empty(9, 32, 64, dtype=torch.bfloat16, device="cuda")
    wft = torch.empty(1, 128, 64, dtype=torch.bfloat16, device="cuda")
    wdt = torch.empty(4, 64, 32, dtype=torch.bfloat16, device="cuda")
    b4 = torch.empty(128, device="cuda")
    wf1 = torch.full((9, 32, 16), 7.0, dtype=torch.bfloat16, device="cuda")
    w3b = rand_f32(128, 96, 3, 3, seed=10)  # several 32x32 tiles in both directions (shared-memory tiled path)
    wf3b = torch.empty(9, 128, 96, dtype=torch.bfloat16, device="cuda")
    wd3b = torch.empty(9, 96, 128, dtype=torch.bfloat16, device="cuda")
    e = [N.PrepEntry(w3.data_ptr(), None, wf3.data_ptr(), wd3.data_ptr(), None, 64, 32, 32, 0),
         N.PrepEntry(wt.data_ptr(), bt.data_ptr(), wft.data_ptr(), wdt.data_ptr(), b4.data_ptr(), 32, 64, 64, 1),
         N.PrepEntry(w1.data_ptr(), None, wf1.data_ptr(), None, None, 32, 3, 16, 0),
         N.PrepEntry(w3b.data_ptr(), None, wf3b.data_ptr(), wd3b.data_ptr(), None, 128, 96, 96, 0)]
    table = N.prep_weights(e, "cuda")
    N.run_prep_weights(table, len(e))
    torch.cuda.synchronize()
    assert torch.equal(wf3, R.conv3x3_weight_fwd(w3))
    assert torch.equal(wd3, R.conv3x3_weight_bwd(w3))
    assert torch.equal(wft, R.convT_weight_fwd(wt))
    assert torch.equal(wdt, R.convT_weight_bwd(wt))
    assert torch.equal(b4, bt.repeat(4))
    assert torch.equal(wf1, R.conv3x3_weight_fwd(w1, 16))
    assert torch.equal(wf3b, R.conv3x3_weight_fwd(w3b))
    assert torch.equal(wd3b, R.conv3x3_weight_bwd(w3b))


def _head_buffers(P, c):
    blocks = N.lib().dpt_head_blocks()
    return (torch.empty(blocks, 4, device="cuda"), torch.empty(blocks, c + 1, device="cuda"),
            torch.zeros(8, device="cuda"), blocks)


@pytest.mark.parametrize("target_dtype", [0, 1])
def test_head_sigmoid_loss_forward_backward(target_dtype):
    B, H, W, c = 2, 40, 60, 32
    P = B * H * W
    x = torch.relu(rand_bf16(B, H, W, c, seed=10).float()).to(torch.bfloat16)
    w = rand_f32(c, scale=0.3, seed=11)
    b = rand_f32(1, scale=0.1, seed=12)
    t = (torch.rand(B, H, W, device="cuda") > 0.6)
    tgt = t.to(torch.uint8) if target_dtype == 1 else t.float()
    prob = torch.empty(B, H, W, device="cuda")
    partial, partial_w, state, blocks = _head_buffers(P, c)
    L = N.lib()
    s = N.stream_ptr()
    N.check(L.dpt_head_sigmoid_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), tgt.data_ptr(), target_dtype,
                                   prob.data_ptr(), partial.data_ptr(), P, c, s))
    parts = (C.c_void_p * 1)(partial.data_ptr())
    nb = (C.c_int * 1)(blocks)
    N.check(L.dpt_head_finalize(parts, nb, 1, float(P), 1.0, state.data_ptr(), None, 0, s))
    dx = torch.empty_like(x)
    dw = torch.empty(c, device="cuda")
    db = torch.empty(1, device="cuda")
    gscale = 4.0
    N.check(L.dpt_head_sigmoid_bwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), tgt.data_ptr(), target_dtype,
                                   state.data_ptr(), dx.data_ptr(), partial_w.data_ptr(), dw.data_ptr(), db.data_ptr(),
                                   P, c, gscale, 1.0, 1, s))
    torch.cuda.synchronize()
    # oracle: the reference's own formulation on probabilities, fp32 autograd
    xr = x.float().requires_grad_(True)
    wr = w.clone().requires_grad_(True)
    br = b.clone().requires_grad_(True)
    p_ref = torch.sigmoid(xr @ wr + br)
    loss_ref = R.bce_dice_loss(p_ref, t.float(), 1.0)
    (gscale * loss_ref).backward()
    assert_close(prob, p_ref.detach(), 1e-5, 1e-4, "prob")
    assert_close(state[4], loss_ref.detach(), 1e-4, 1e-4, "loss")
    assert_close(dw, wr.grad, 1e-4, 2e-3, "head dw")
    assert_close(db, br.grad, 1e-4, 2e-3, "head db")
    assert_close(dx, xr.grad * (x.float() > 0), 1e-6, 2e-2, "head dx")


def test_adam_matches_torch():
    n = 1000003 // 4 * 4 + 2  # exercises the scalar tail
    p = rand_f32(n, seed=13)
    pr = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=1e-3, weight_decay=1e-8)
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    state = torch.zeros(8, device="cuda")
    L = N.lib()
    s = N.stream_ptr()
    N.check(L.dpt_opt_init(state.data_ptr(), 1e-3, 0, 0.9, 0.999, s))
    for it in range(3):
        g = rand_f32(n, seed=20 + it)
        pr.grad = g.clone()
        opt.step()
        N.check(L.dpt_opt_tick(state.data_ptr(), s))
        N.check(L.dpt_adam_step(p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), n, state.data_ptr(), 1e-8, 1e-8,
                                1.0, 1, s))
        torch.cuda.synchronize()
        assert float(g.abs().max()) == 0.0, "zero_grad fusion"
    assert_close(p, pr.detach(), 1e-6, 1e-5, "adam params")


def test_rmsprop_matches_torch():
    n = 4096 * 3 + 1
    p = rand_f32(n, seed=14)
    pr = p.clone().requires_grad_(True)
    opt = torch.optim.RMSprop([pr], lr=1e-3, weight_decay=1e-8, momentum=0.9)
    sq = torch.zeros(n, device="cuda")
    buf = torch.zeros(n, device="cuda")
    state = torch.zeros(8, device="cuda")
    L = N.lib()
    s = N.stream_ptr()
    N.check(L.dpt_opt_init(state.data_ptr(), 1e-3, 0, 0.9, 0.999, s))
    for it in range(3):
        g = rand_f32(n, seed=30 + it)
        pr.grad = g.clone()
        opt.step()
        N.check(L.dpt_rmsprop_step(p.data_ptr(), g.data_ptr(), sq.data_ptr(), buf.data_ptr(), n, state.data_ptr(), 0.99,
                                   1e-8, 1e-8, 0.9, 1.0, 0, s))
    torch.cuda.synchronize()
    assert_close(p, pr.detach(), 1e-6, 1e-5, "rmsprop params")


def test_single_rank_team_collectives():
    # world = 1 exercises the barrier protocol, the fused optimizer epilogue and the epoch bookkeeping.
    n = 4096 * 5
    L = N.lib()
    s = N.stream_ptr()
    pad = torch.zeros(N.PAD_BYTES // 4, dtype=torch.int32, device="cuda")
    epoch = torch.zeros(N.EPOCH_BYTES // 4, dtype=torch.int32, device="cuda")
    team = N.Team()
    team.rank, team.world = 0, 1
    team.pads[0] = pad.data_ptr()
    team.epoch = epoch.data_ptr()
    p = rand_f32(n, seed=15)
    pr = p.clone().requires_grad_(True)
    opt = torch.optim.Adam([pr], lr=1e-3, weight_decay=1e-8)
    m = torch.zeros(n, device="cuda")
    v = torch.zeros(n, device="cuda")
    state = torch.zeros(8, device="cuda")
    N.check(L.dpt_opt_init(state.data_ptr(), 1e-3, 0, 0.9, 0.999, s))
    for algo in (N.AR_TWO_SHOT, N.AR_ONE_SHOT):
        g = rand_f32(n, seed=40 + algo)
        pr.grad = 0.5 * g
        opt.step()
        d = N.FusedArDesc()
        d.grads[0] = g.data_ptr()
        d.params[0] = p.data_ptr()
        d.state0, d.state1, d.opt_state = m.data_ptr(), v.data_ptr(), state.data_ptr()
        d.n, d.algo, d.optimizer, d.blocks = n, algo, N.OPT_ADAM, 16
        d.eps, d.weight_decay, d.grad_scale = 1e-8, 1e-8, 0.5
        N.check(L.dpt_opt_tick(state.data_ptr(), s))
        N.check(L.dpt_fused_allreduce_opt(C.byref(team), C.byref(d), s))
    torch.cuda.synchronize()
    assert_close(p, pr.detach(), 1e-6, 1e-5, "fused allreduce+adam (world=1)")
    bufs = (C.c_void_p * 1)(p.data_ptr())
    before = p.clone()
    N.check(L.dpt_team_allreduce_sum(C.byref(team), bufs, n, 2.0, 8, s))
    torch.cuda.synchronize()
    assert_close(p, 2 * before, 1e-6, 1e-6, "allreduce_sum (world=1)")
