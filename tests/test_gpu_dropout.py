"""Dropout > 0 in training mode (reference: nn.MultiheadAttention(dropout=p) on the attention
probabilities, Residual's nn.Dropout on the branch output).  The kernels draw their masks from a
counter-based hash, not from torch's Philox stream, so the checks are: same sites / scaling / drop rate,
reproducibility under torch.manual_seed, identity in eval mode, and gradients that differentiate exactly
the masked forward (finite differences with the mask held fixed; bf16 against fp32 with the same seed)."""
import pytest
import torch

from gpu_util import assert_rel, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


def test_residual_dropout_rate_and_scaling():
    from perceiver import Residual, mlp
    torch.manual_seed(0)
    p = 0.25
    res = Residual(mlp(64), p).to(DEV)
    x = torch.randn(64, 32, 64, device=DEV)
    res.eval()
    y0 = res(x)                      # eval mode: dropout is the identity
    res.train()
    y = res(x)
    branch0, branch = y0 - x, y - x
    dropped = branch == 0
    frac = dropped.float().mean().item()
    assert abs(frac - p) < 0.01, frac
    kept = ~dropped
    assert_rel(branch[kept], branch0[kept] / (1 - p), 1e-5, "kept values are scaled by 1/(1-p)")


def test_dropout_is_reproducible_under_manual_seed_and_off_in_eval():
    from perceiver import PerceiverEncoder, TextInputAdapter
    import perceiver.functional as F
    enc = PerceiverEncoder(TextInputAdapter(50, 16, 32), (8, 32), 2, 2, 2, 2, dropout=0.1).to(DEV)
    ids = torch.randint(3, 50, (4, 16), device=DEV)

    def run():
        torch.manual_seed(123)
        F._drop_calls[0] = 0
        return enc(ids)[0]
    a, b = run(), run()
    assert torch.equal(a, b)
    F._drop_calls[0] = 77
    c = enc(ids)[0]
    assert rel_err(c, a) > 1e-3       # another seed, another mask
    enc.eval()
    e1, e2 = enc(ids)[0], enc(ids)[0]
    assert torch.equal(e1, e2) and rel_err(e1, a) > 1e-3


def _fd_check(fn, inputs, idx_list, eps=1e-2, tol=3e-2):
    """central finite differences of sum(fn(*inputs) * w) at a few coordinates vs autograd"""
    torch.manual_seed(5)
    out = fn(*inputs)
    w = torch.randn_like(out)
    (out * w).sum().backward()
    for ti, coords in idx_list:
        t = inputs[ti]
        for c in coords:
            orig = t.data[c].item()
            t.data[c] = orig + eps
            up = (fn(*inputs) * w).sum().item()
            t.data[c] = orig - eps
            dn = (fn(*inputs) * w).sum().item()
            t.data[c] = orig
            fd = (up - dn) / (2 * eps)
            an = t.grad[c].item()
            assert abs(fd - an) <= tol * max(1.0, abs(fd), abs(an)), (ti, c, fd, an)


def test_self_attention_layer_gradients_differentiate_the_masked_forward():
    import perceiver.functional as F
    torch.manual_seed(1)
    b, n, c, heads, p, seed = 2, 6, 16, 2, 0.3, 987654321
    x = torch.randn(b, n, c, device=DEV, requires_grad=True)
    ln_w = (torch.randn(c, device=DEV) * 0.2 + 1).requires_grad_(True)
    ln_b = (torch.randn(c, device=DEV) * 0.1).requires_grad_(True)
    in_w = (torch.randn(3 * c, c, device=DEV) * 0.3).requires_grad_(True)
    in_b = (torch.randn(3 * c, device=DEV) * 0.1).requires_grad_(True)
    wo = (torch.randn(c, c, device=DEV) * 0.3).requires_grad_(True)
    bo = (torch.randn(c, device=DEV) * 0.1).requires_grad_(True)
    inputs = [x, ln_w, ln_b, in_w, in_b, wo, bo]

    def fn(*t):
        return F.SelfAttnFn.apply(*t, heads, True, p, seed)
    _fd_check(fn, inputs, [(0, [(0, 0, 0), (1, 3, 7), (0, 5, 15)]), (3, [(0, 0), (20, 5), (40, 9)]), (5, [(3, 3), (10, 1)]),
                           (1, [(2,)]), (6, [(4,)])])


def test_cross_attention_and_mlp_gradients_with_dropout():
    import perceiver.functional as F
    torch.manual_seed(2)
    b, n, m, e, heads, p, seed = 2, 5, 9, 16, 2, 0.2, 424242
    x_q = torch.randn(b, n, e, device=DEV, requires_grad=True)
    kv = torch.randn(b, m, 2 * e, device=DEV, requires_grad=True)
    pad = torch.zeros(b, m, dtype=torch.bool, device=DEV)
    pad[1, 6:] = True
    ln_w = (torch.randn(e, device=DEV) * 0.2 + 1).requires_grad_(True)
    ln_b = (torch.randn(e, device=DEV) * 0.1).requires_grad_(True)
    bias = (torch.randn(3 * e, device=DEV) * 0.1).requires_grad_(True)
    w_packed = (torch.randn(3 * e, e, device=DEV) * 0.3).requires_grad_(True)
    wo = (torch.randn(e, e, device=DEV) * 0.3).requires_grad_(True)
    bo = (torch.randn(e, device=DEV) * 0.1).requires_grad_(True)

    def fn(x_q, kv, ln_w, ln_b, bias, w_packed, wo, bo):
        return F.CrossAttendFn.apply(x_q, kv, pad, ln_w, ln_b, bias, w_packed, None, wo, bo, heads, True, p, seed)
    _fd_check(fn, [x_q, kv, ln_w, ln_b, bias, w_packed, wo, bo],
              [(0, [(0, 0, 0), (1, 4, 9)]), (1, [(0, 0, 0), (1, 2, 20), (0, 8, 31)]), (5, [(0, 0), (7, 3)]), (6, [(1, 1)])])
    x = torch.randn(3, 4, e, device=DEV, requires_grad=True)
    ps = [(torch.randn(s, device=DEV) * 0.3).requires_grad_(True) for s in [(e,), (e,), (e, e), (e,), (e, e), (e,)]]

    def mfn(x, *ps):
        return F.MlpFn.apply(x, *ps, True, p, seed)
    _fd_check(mfn, [x, *ps], [(0, [(0, 0, 0), (2, 3, 5)]), (3, [(1, 2)]), (5, [(4, 4)]), (6, [(3,)])])


def test_bf16_mode_with_dropout_matches_fp32_mode_with_the_same_masks():
    """the mask is a function of (seed, index) only, so both arithmetic modes drop the same elements"""
    import perceiver
    import perceiver.functional as F
    from perceiver import PerceiverEncoder, TextInputAdapter
    torch.manual_seed(3)
    enc = PerceiverEncoder(TextInputAdapter(100, 64, 128), (32, 128), 2, 2, 2, 2, dropout=0.1).to(DEV)
    ids = torch.randint(3, 100, (4, 64), device=DEV)
    outs, grads = {}, {}
    try:
        for mode in ("fp32", "bf16"):
            perceiver.set_compute_mode(mode)
            enc.zero_grad(set_to_none=True)
            torch.manual_seed(9)
            F._drop_calls[0] = 0
            lat, _ = enc(ids)
            lat.square().mean().backward()
            outs[mode] = lat.detach().clone()
            grads[mode] = {k: p.grad.detach().clone() for k, p in enc.named_parameters()}
    finally:
        perceiver.set_compute_mode("fp32")
    assert_rel(outs["bf16"], outs["fp32"], 2e-2, "latents")
    num = sum(((grads["bf16"][k] - grads["fp32"][k]).double().norm() ** 2).item() for k in grads["fp32"])
    den = sum((grads["fp32"][k].double().norm() ** 2).item() for k in grads["fp32"])
    assert (num / den) ** 0.5 < 2e-2


def test_masks_of_neighbouring_sites_are_uncorrelated():
    """with hash(idx + seed) the mask of site seed + k was the mask of site seed shifted by k elements (the
    Residual site is seed + 1, latent layer s uses seed + 4 s and + 2).  The site key is now mixed on its own before it
    meets the element index: masks of neighbouring seeds must be uncorrelated at every small shift."""
    import perceiver.functional as F
    torch.manual_seed(0)
    p, n, c = 0.5, 4096, 64
    ps = [torch.ones(c, device=DEV), torch.zeros(c, device=DEV), torch.eye(c, device=DEV), torch.zeros(c, device=DEV),
          torch.eye(c, device=DEV), torch.zeros(c, device=DEV)]
    x = torch.randn(n, c, device=DEV) + 3.0

    def mask(seed):
        # Residual(mlp) with dropout: y - x is zero exactly where the branch output was dropped (site seed + 1)
        y = F.MlpFn.apply(x, *ps, True, p, seed)
        return ((y - x) != 0).float().reshape(-1)

    base = 123456789
    m0 = mask(base)
    assert abs(m0.mean().item() - (1 - p)) < 0.01
    for dseed in (1, 2, 3, 4, 8):
        m1 = mask(base + dseed)
        for shift in range(0, 9):
            a = m0[: m0.numel() - 16]
            b = m1[shift: shift + a.numel()]
            corr = ((a - a.mean()) * (b - b.mean())).mean() / (a.std() * b.std())
            assert abs(corr.item()) < 0.02, (dseed, shift, corr.item())
