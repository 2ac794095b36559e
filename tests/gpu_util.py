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


def grad_table(got: dict, ref: dict, min_abs=1e-10):
    """per-tensor (name, max-abs error / max |ref|, relative L2, numel) sorted worst first, and the relative L2 error of
    the whole gradient (all tensors concatenated)"""
    rows, num, den = [], 0.0, 0.0
    for k, r in ref.items():
        if r is None or r.abs().max() < min_abs:
            continue
        g = got[k].detach().double().cpu()
        r = r.detach().double().cpu()
        d = (g - r).norm().item()
        rows.append((k, (g - r).abs().max().item() / r.abs().max().item(), d / max(r.norm().item(), 1e-300), r.numel()))
        num += d * d
        den += r.norm().item() ** 2
    rows.sort(key=lambda t: -t[1])
    return rows, (num / max(den, 1e-300)) ** 0.5


# bf16-mode per-tensor gradient bounds (max-abs error over the tensor's largest magnitude; note that this, like
# rel_err above, is a tensor-scale metric, not an element-wise relative error).  The north-star tolerance 1e-2 holds
# for the gradient as a whole (relative L2 over all parameters: measured 7.0e-3 at the benchmarked cfg2 batch 64).
# Per tensor, measured against the fp32 oracle on B200 (tests/test_gpu_train_step.py -s, cfg2 batch 64 and the small
# config): every matrix <= 1.3e-2, every LayerNorm / bias vector <= 2.2e-2, the LayerNorm affine terms of the decoder
# queries <= 3.3e-2 — bounded here by 2e-2, 3e-2 and 5e-2.
# (Round 1 held the decoder-query terms to 1e-1 and blamed cancellation.  The cause was the attention backward's row
#  term delta = dO.O: taken from a rounded O it is inconsistent with the recomputed probabilities, sum_k dS_k is then
#  -error instead of 0, and dQ gets a component along the sample's mean key that is common to all rows.  The kernels
#  that see the whole key axis in one tile now form delta = sum_k p_k dP_k from the recomputed p — pio_attn_tc.cu dQ
#  kernel for <= 64 keys, the labelled-row decoder kernel — and the error fell from 1.3e-1 to 2.5e-2.)
BF16_GRAD_BOUNDS = [
    ("decoder.cross_attention.0.module.q_norm.", 5e-2), ("1.cross_attention.0.module.q_norm.", 5e-2),
    ("norm.", 3e-2), (".module.0.", 3e-2), ("bias", 3e-2), ("latent", 3e-2), ("output", 3e-2),
    ("", 2e-2),
]


def bf16_grad_bound(name: str) -> float:
    for frag, tol in BF16_GRAD_BOUNDS:
        if frag in name:
            return tol
    return 2e-2
