import torch

FP32_TOL = 1e-4   # north_star: fp32 mode, rel 1e-4
BF16_TOL = 1e-2   # north_star: bf16 mode, rel 1e-2


def rel_err(got: torch.Tensor, ref: torch.Tensor) -> float:
    """max |got-ref| relative to the largest reference magnitude (scale-aware, robust to zeros)."""
    got = got.detach().double().cpu()
    ref = ref.detach().double().cpu()
    assert got.shape == ref.shape, (got.shape, ref.shape)
    if ref.numel() == 0:
        return 0.0
    denom = max(ref.abs().max().item(), 1e-12)
    return (got - ref).abs().max().item() / denom


def assert_rel(got, ref, tol, what=""):
    assert torch.isfinite(got).all(), f"{what}: non-finite values"
    e = rel_err(got, ref)
    assert e <= tol, f"{what}: rel err {e:.3e} > {tol:.1e}"
