"""Helpers shared by the GPU numerics tests."""
import torch


def rand_bf16(*shape, scale=1.0, device="cuda", seed=None):
    g = torch.Generator(device=device)
    if seed is not None:
        g.manual_seed(seed)
    else:
        g.manual_seed(abs(hash(shape)) % (2**31))
    return ((torch.rand(*shape, device=device, generator=g) * 2 - 1) * scale).to(torch.bfloat16)


def rand_f32(*shape, scale=1.0, device="cuda", seed=0):
    g = torch.Generator(device=device)
    g.manual_seed(seed)
    return (torch.rand(*shape, device=device, generator=g) * 2 - 1) * scale


def assert_close(got, ref, atol, rtol, what=""):
    got = got.float()
    ref = ref.float()
    assert got.shape == ref.shape, f"{what}: shape {tuple(got.shape)} vs {tuple(ref.shape)}"
    assert torch.isfinite(got).all(), f"{what}: non-finite values in result"
    err = (got - ref).abs()
    tol = atol + rtol * ref.abs()
    bad = err > tol
    if bad.any():
        idx = torch.nonzero(bad)[0].tolist()
        raise AssertionError(
            f"{what}: {int(bad.sum())}/{bad.numel()} elements off; max abs err {err.max().item():.4g} "
            f"(ref max {ref.abs().max().item():.4g}); first bad index {idx}: got {got[tuple(idx)].item():.5g} "
            f"ref {ref[tuple(idx)].item():.5g}")
