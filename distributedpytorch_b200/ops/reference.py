"""Plain-PyTorch fp32 oracles for every native kernel (same op, same layouts), used by the GPU numerics
tests and by the CPU plumbing path.  Activations are NHWC; weights are PyTorch-layout fp32 tensors."""
from __future__ import annotations

from typing import Optional, Sequence

import torch
import torch.nn.functional as F


def nhwc_to_nchw(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 3, 1, 2).contiguous()


def nchw_to_nhwc(x: torch.Tensor) -> torch.Tensor:
    return x.permute(0, 2, 3, 1).contiguous()


# ---- kernel-side weight layouts (mirror csrc/kernels/elementwise.cu:prep_weights_kernel) ------------------

def conv3x3_weight_fwd(w: torch.Tensor, cin_pad: Optional[int] = None) -> torch.Tensor:
    """[Cout,Cin,3,3] fp32 -> bf16 [9][Cout][Cin_pad]"""
    cout, cin = w.shape[:2]
    cin_pad = cin_pad or cin
    out = torch.zeros(9, cout, cin_pad, dtype=torch.bfloat16, device=w.device)
    out[:, :, :cin] = w.reshape(cout, cin, 9).permute(2, 0, 1).to(torch.bfloat16)
    return out.contiguous()


def conv3x3_weight_bwd(w: torch.Tensor) -> torch.Tensor:
    """[Cout,Cin,3,3] fp32 -> bf16 [9][Cin][Cout] with the taps flipped (backward-data as a forward conv)"""
    cout, cin = w.shape[:2]
    return w.reshape(cout, cin, 9).flip(2).permute(2, 1, 0).to(torch.bfloat16).contiguous()


def convT_weight_fwd(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d [Cin,Cout,2,2] fp32 -> bf16 [1][4*Cout][Cin] (row = q*Cout+co, q = dy*2+dx)"""
    cin, cout = w.shape[:2]
    return w.reshape(cin, cout, 4).permute(2, 1, 0).reshape(1, 4 * cout, cin).to(torch.bfloat16).contiguous()


def convT_weight_bwd(w: torch.Tensor) -> torch.Tensor:
    """ConvTranspose2d [Cin,Cout,2,2] fp32 -> bf16 [4][Cin][Cout]"""
    cin, cout = w.shape[:2]
    return w.reshape(cin, cout, 4).permute(2, 0, 1).to(torch.bfloat16).contiguous()


# ---- ops ------------------------------------------------------------------------------------------------

def conv3x3(srcs: Sequence[torch.Tensor], w: torch.Tensor, bias: Optional[torch.Tensor], relu: bool) -> torch.Tensor:
    """srcs: NHWC tensors concatenated along C; w fp32 [Cout, Cin_total(<= sum C, rest is padding), 3, 3]."""
    x = torch.cat([s.float() for s in srcs], dim=-1)
    cin = w.shape[1]
    y = F.conv2d(nhwc_to_nchw(x[..., :cin]), w.float(), bias.float() if bias is not None else None, padding=1)
    if relu:
        y = F.relu(y)
    return nchw_to_nhwc(y)


def conv3x3_dgrad(dz: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    """dz NHWC [B,H,W,Cout], w [Cout,Cin,3,3] -> dx NHWC [B,H,W,Cin]"""
    return nchw_to_nhwc(F.conv_transpose2d(nhwc_to_nchw(dz.float()), w.float(), padding=1))


def conv3x3_wgrad(x: torch.Tensor, dz: torch.Tensor) -> torch.Tensor:
    """-> [Cout,Cin,3,3]"""
    xn, dn = nhwc_to_nchw(x.float()), nhwc_to_nchw(dz.float())
    cin, cout = xn.shape[1], dn.shape[1]
    return torch.nn.grad.conv2d_weight(xn, (cout, cin, 3, 3), dn, padding=1)


def convT2x2(x: torch.Tensor, w: torch.Tensor, bias: Optional[torch.Tensor]) -> torch.Tensor:
    """x NHWC [B,H,W,Cin], w [Cin,Cout,2,2] -> NHWC [B,2H,2W,Cout]"""
    return nchw_to_nhwc(F.conv_transpose2d(nhwc_to_nchw(x.float()), w.float(),
                                           bias.float() if bias is not None else None, stride=2))


def convT2x2_dgrad(dy: torch.Tensor, w: torch.Tensor) -> torch.Tensor:
    return nchw_to_nhwc(F.conv2d(nhwc_to_nchw(dy.float()), w.float(), stride=2))


def convT2x2_wgrad(x: torch.Tensor, dy: torch.Tensor) -> torch.Tensor:
    """-> [Cin,Cout,2,2]"""
    xn, dn = nhwc_to_nchw(x.float()), nhwc_to_nchw(dy.float())
    # conv_transpose2d(x, w) == gradient of conv2d wrt its input; its weight gradient is conv2d_weight with
    # the roles of input and grad_output exchanged.
    return torch.nn.grad.conv2d_weight(dn, (xn.shape[1], dn.shape[1], 2, 2), xn, stride=2)


def maxpool2x2(x: torch.Tensor) -> torch.Tensor:
    return nchw_to_nhwc(F.max_pool2d(nhwc_to_nchw(x.float()), 2, 2))


def pool_bwd_fused(act: torch.Tensor, d_pool: torch.Tensor, d_skip: Optional[torch.Tensor], relu_mask: bool):
    a = nhwc_to_nchw(act.float()).requires_grad_(True)
    p = F.max_pool2d(a, 2, 2)
    (g,) = torch.autograd.grad(p, a, nhwc_to_nchw(d_pool.float()))
    g = nchw_to_nhwc(g)
    if d_skip is not None:
        g = g + d_skip.float()
    if relu_mask:
        g = g * (act.float() > 0)
    return g


def bce_dice_loss(prob: torch.Tensor, target: torch.Tensor, dice_weight: float = 1.0) -> torch.Tensor:
    """utils/utils.py:9-25 of the reference in closed form (BCE mean - log(2I/U), Dice over the whole batch)."""
    loss = F.binary_cross_entropy(prob, target)
    if dice_weight:
        t1 = (target == 1).float()
        inter = (prob * t1).sum()
        union = prob.sum() + t1.sum() + 1e-15
        loss = loss - dice_weight * torch.log(2 * inter / union)
    return loss
