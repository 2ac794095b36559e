"""The fused latent-block kernels (one launch per self_attention_block, reference model.py:36-44) against (a) the
unfused bf16 operators they replace and (b) the fp32 oracle: the cluster kernel (pio_latent_cl.cu, four CTAs per
128-row tile, the default up to 66 tiles) and the one-CTA-per-tile kernel (pio_latent_tc.cu).  The fused forward
writes the same saved-for-backward state as the unfused one, so gradients through it check every saved tensor.  The
backward runs through the fused input-gradient chain (pio_latent_bwd_cl.cu + the batched parameter-gradient kernels)
or through the unfused operators."""
import os

import pytest
import torch

from gpu_util import BF16_TOL, assert_rel, rel_err
from oracle import perceiver_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True, params=["cluster", "single", "cluster-fwd-only"])
def bf16_mode(request):
    import perceiver
    perceiver.set_compute_mode("bf16")
    os.environ.pop("PIO_NO_FUSED_LATENT", None)
    os.environ["PIO_LATENT_CLUSTER"] = "0" if request.param == "single" else "1"  # (read per call)
    os.environ["PIO_LATENT_BWD_CLUSTER"] = "1" if request.param == "cluster" else "0"
    os.environ["PIO_FUSED_LATENT_MIN_ROWS"] = "0"  # by default the one-CTA kernel only takes blocks of >= 56 row tiles
    yield
    os.environ.pop("PIO_NO_FUSED_LATENT", None)
    os.environ.pop("PIO_FUSED_LATENT_MIN_ROWS", None)
    os.environ.pop("PIO_LATENT_CLUSTER", None)
    os.environ.pop("PIO_LATENT_BWD_CLUSTER", None)
    perceiver.set_compute_mode("fp32")


def _run(block, x, dy):
    block.zero_grad(set_to_none=True)
    xin = x.clone().requires_grad_(True)
    y = block(xin)
    y.backward(dy)
    torch.cuda.synchronize()
    return y.detach(), xin.grad.detach(), {k: p.grad.detach().clone() for k, p in block.named_parameters()}


@pytest.mark.parametrize("batch,n,layers", [(2, 64, 1), (4, 64, 2), (3, 64, 6), (5, 32, 3), (3, 128, 2), (2, 64, 9), (64, 64, 6), (128, 64, 2)])
def test_fused_block_matches_unfused_operators(batch, n, layers):
    from perceiver import _native as N
    from perceiver.model import self_attention_block
    torch.manual_seed(batch * 100 + n + layers)
    block = self_attention_block(layers, 256, 4, 0.0).to(DEV)
    with torch.no_grad():  # LayerNorm affine terms and biases away from their 1 / 0 initial values
        for k, p in block.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x = torch.randn(batch, n, 256, device=DEV)
    dy = torch.randn(batch, n, 256, device=DEV)
    c0 = N.lib().pio_launch_count()
    y_f, dx_f, g_f = _run(block, x, dy)
    fused_launches = N.lib().pio_launch_count() - c0
    os.environ["PIO_NO_FUSED_LATENT"] = "1"
    c0 = N.lib().pio_launch_count()
    y_u, dx_u, g_u = _run(block, x, dy)
    unfused_launches = N.lib().pio_launch_count() - c0
    os.environ.pop("PIO_NO_FUSED_LATENT")
    assert fused_launches < unfused_launches - 5 * layers, (fused_launches, unfused_launches)  # the fused path really ran
    assert_rel(y_f, y_u, 4e-3, "y (fused vs unfused)")
    assert_rel(dx_f, dx_u, 1e-2, "dx (fused vs unfused)")
    for k in g_u:
        assert_rel(g_f[k], g_u[k], 1.5e-2, f"grad {k} (fused vs unfused)")


@pytest.mark.parametrize("batch,n,layers", [(4, 64, 3), (6, 32, 2)])
def test_fused_block_matches_oracle(batch, n, layers):
    from perceiver.model import self_attention_block
    torch.manual_seed(3)
    block = self_attention_block(layers, 256, 4, 0.0)
    with torch.no_grad():
        for k, p in block.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in block.state_dict().items()}
    x = torch.randn(batch, n, 256)
    xr = x.clone().requires_grad_(True)
    ref = xr
    for s in range(layers):
        ref = O.self_attention_residual(sd, f"{s}.0.", ref, 4)
        ref = O.mlp_residual(sd, f"{s}.1.", ref)
    dy = torch.randn_like(ref)
    ref.backward(dy)
    block = block.to(DEV)
    y, dx, g = _run(block, x.to(DEV), dy.to(DEV))
    assert_rel(y, ref, BF16_TOL, "y")
    assert_rel(dx, xr.grad, 2e-2, "dx")
    num = sum(((g[k].cpu().double() - sd[k].grad.double()) ** 2).sum().item() for k in g)
    den = sum((sd[k].grad.double() ** 2).sum().item() for k in g)
    assert (num / den) ** 0.5 < BF16_TOL, (num / den) ** 0.5


@pytest.mark.parametrize("batch,n,layers", [(2, 64, 1), (5, 32, 3), (64, 64, 6), (3, 128, 7)])
def test_pre_mlp_rides_on_the_block(batch, n, layers):
    """Residual(mlp) of the cross-attention layer handed to the block's fused call as an MLP-only first layer
    (reference model.py:29-33 followed by :36-44) against running it as an operator of its own."""
    from perceiver import _native as N
    from perceiver.model import Residual, mlp, self_attention_block
    torch.manual_seed(11 * batch + n + layers)
    block = self_attention_block(layers, 256, 4, 0.0).to(DEV)
    pre = Residual(mlp(256), 0.0).to(DEV)
    with torch.no_grad():
        for mod in (block, pre):
            for k, p in mod.named_parameters():
                if p.dim() == 1:
                    p.add_(torch.randn_like(p) * 0.1)
    x = torch.randn(batch, n, 256, device=DEV)
    dy = torch.randn(batch, n, 256, device=DEV)

    def run(fused):
        block.zero_grad(set_to_none=True)
        pre.zero_grad(set_to_none=True)
        xin = x.clone().requires_grad_(True)
        c0 = N.lib().pio_launch_count()
        y = block(xin, pre=pre) if fused else block(pre(xin))
        y.backward(dy)
        torch.cuda.synchronize()
        launches = N.lib().pio_launch_count() - c0
        grads = {"block." + k: p.grad.detach().clone() for k, p in block.named_parameters()}
        grads.update({"pre." + k: p.grad.detach().clone() for k, p in pre.named_parameters()})
        return y.detach(), xin.grad.detach(), grads, launches

    y_f, dx_f, g_f, l_f = run(True)
    y_u, dx_u, g_u, l_u = run(False)
    if os.environ["PIO_LATENT_CLUSTER"] == "1" and os.environ["PIO_LATENT_BWD_CLUSTER"] == "1":
        assert l_f < l_u, (l_f, l_u)  # the MLP's own launches are gone
    assert_rel(y_f, y_u, 4e-3, "y")
    assert_rel(dx_f, dx_u, 1e-2, "dx")
    for k in g_u:
        assert_rel(g_f[k], g_u[k], 1.5e-2, f"grad {k}")


@pytest.mark.parametrize("batch,n,m,layers,heads_x", [(2, 64, 96, 1, 4), (5, 32, 50, 2, 2), (64, 64, 512, 6, 4)])
def test_cross_layer_out_projection_rides_on_the_block(batch, n, m, layers, heads_x):
    """A whole encoder layer (cross_attention_layer + self_attention_block, reference model.py:29-44, :186-191) with the
    cross-attention layer's out-projection, residual and MLP handed to the block's fused call, against the layer run
    operator by operator."""
    from perceiver.model import PerceiverEncoder, cross_attention_layer, self_attention_block
    from perceiver.utils import Sequential
    torch.manual_seed(7 * batch + n + m + layers)
    layer = Sequential(cross_attention_layer(256, 256, heads_x, 0.0), self_attention_block(layers, 256, 4, 0.0)).to(DEV)
    with torch.no_grad():
        for k, p in layer.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x_lat = torch.randn(batch, n, 256, device=DEV)
    x_in = torch.randn(batch, m, 256, device=DEV)
    pad = torch.zeros(batch, m, dtype=torch.bool, device=DEV)
    pad[:, m - 5:] = True
    dy = torch.randn(batch, n, 256, device=DEV)

    def run(fused):
        if fused:
            os.environ.pop("PIO_NO_PRE_PROJ", None)
            os.environ.pop("PIO_NO_PRE_MLP", None)
        else:
            os.environ["PIO_NO_PRE_PROJ"] = "1"
            os.environ["PIO_NO_PRE_MLP"] = "1"
        layer.zero_grad(set_to_none=True)
        a = x_lat.clone().requires_grad_(True)
        b = x_in.clone().requires_grad_(True)
        y = PerceiverEncoder._apply_layer(layer, a, b, pad, None)
        y.backward(dy)
        torch.cuda.synchronize()
        return y.detach(), a.grad.detach(), b.grad.detach(), {k: p.grad.detach().clone() for k, p in layer.named_parameters()}

    try:
        y_f, da_f, db_f, g_f = run(True)
        y_u, da_u, db_u, g_u = run(False)
    finally:
        os.environ.pop("PIO_NO_PRE_PROJ", None)
        os.environ.pop("PIO_NO_PRE_MLP", None)
    assert_rel(y_f, y_u, 4e-3, "y")
    assert_rel(da_f, da_u, 1e-2, "d latent")
    assert_rel(db_f, db_u, 1.5e-2, "d input")
    for k in g_u:
        assert_rel(g_f[k], g_u[k], 1.5e-2, f"grad {k}")
