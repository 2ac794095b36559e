"""The decoder evaluated for the labelled (sample, position) pairs only (pio_label_compact + pio_decode_rows_fwd/bwd,
reference model.py:229-237 + lightning.py:223-226) against decoding every position: an ignored position contributes
exactly zero to the loss and to every gradient, so loss and gradients must agree."""
import os

import pytest
import torch

from gpu_util import assert_rel, grad_table, rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def bf16_mode():
    import perceiver
    perceiver.set_compute_mode("bf16")
    os.environ.pop("PIO_NO_COMPACT_DECODE", None)
    yield
    os.environ.pop("PIO_NO_COMPACT_DECODE", None)
    perceiver.set_compute_mode("fp32")


@pytest.mark.parametrize("shape,p", [((7, 100), 0.15), ((64, 512), 0.1), ((3, 50), 1.0), ((5, 33), 0.0)])
def test_label_compact_matches_torch(shape, p):
    from perceiver import functional as F
    torch.manual_seed(shape[0])
    classes = 1003
    labels = torch.randint(0, classes, shape, device=DEV)
    labels[torch.rand(shape, device=DEV) >= p] = -100
    idx, labels_c, count = F.label_compact(labels, classes)
    torch.cuda.synchronize()
    want = (labels.reshape(-1) != -100).nonzero().flatten()
    n = int(count.item())
    assert n == want.numel()
    assert torch.equal(idx[:n].long(), want)
    assert torch.equal(labels_c[:n], labels.reshape(-1)[want])
    assert bool((labels_c[n:] == -100).all())


def _mlm(vocab, seq, latent, layers, sa, heads):
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    c = latent[1]
    enc = PerceiverEncoder(TextInputAdapter(vocab, seq, c), latent, layers, heads[0], heads[1], sa)
    dec = PerceiverDecoder(TextOutputAdapter(vocab, seq, c), latent, heads[2])
    return PerceiverMLM(enc, dec, TextMasking(vocab, 1, 2, 3))


@pytest.mark.parametrize("vocab,seq,l,latent,heads,batch,p_label", [
    (1003, 128, 128, (32, 64), (2, 2, 2), 8, 0.15),
    (1003, 128, 100, (32, 64), (2, 2, 4), 5, 0.3),      # l < max_seq_len, 4 decoder heads of 16
    (10003, 512, 512, (64, 256), (4, 4, 4), 4, 0.12),   # cfg2 shapes
    (503, 64, 64, (16, 128), (2, 2, 1), 3, 1.0),        # every position labelled, one decoder head of 128
    (503, 64, 64, (16, 128), (2, 2, 2), 3, 0.02),       # almost nothing labelled (some samples have no labelled row)
])
def test_labelled_row_decoder_matches_full_decoder(vocab, seq, l, latent, heads, batch, p_label):
    from perceiver import _native as N
    torch.manual_seed(0)
    model = _mlm(vocab, seq, latent, 2, 2, heads).to(DEV)
    with torch.no_grad():
        for k, p in model.named_parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.05)
    ids = torch.randint(3, vocab, (batch, l), device=DEV)
    pad = torch.zeros(batch, l, dtype=torch.bool, device=DEV)
    pad[-1, l // 2:] = True
    labels = torch.full_like(ids, -100)
    sel = (torch.rand(ids.shape, device=DEV) < p_label) & ~pad
    sel[0, 0] = True
    labels[sel] = ids[sel]

    def run():
        model.zero_grad(set_to_none=True)
        c0 = N.lib().pio_launch_count()
        loss, _ = model(ids.clone(), pad, masking=False, return_loss=True, labels=labels)
        (loss * 1.7).backward()
        torch.cuda.synchronize()
        return loss.detach().clone(), {k: p.grad.detach().clone() for k, p in model.named_parameters()}, N.lib().pio_launch_count() - c0

    loss_c, g_c, _ = run()
    os.environ["PIO_NO_COMPACT_DECODE"] = "1"
    loss_f, g_f, _ = run()
    os.environ.pop("PIO_NO_COMPACT_DECODE")
    assert torch.isfinite(loss_c)
    assert_rel(loss_c, loss_f, 2e-3, "loss")
    rows, glob = grad_table(g_c, g_f)
    print(f"whole-gradient rel L2 {glob:.3e}; worst: {rows[:4]}")
    assert glob < 1e-2, glob
    for k, e, l2, nel in rows:
        assert e < 5e-2, (k, e)


@pytest.mark.parametrize("batch,l,k_queries,n,e,heads,p_label", [(4, 50, 64, 32, 64, 2, 0.3), (3, 128, 128, 64, 256, 4, 0.15), (2, 40, 40, 16, 128, 1, 1.0)])
def test_decode_rows_operator_against_the_fp64_oracle(batch, l, k_queries, n, e, heads, p_label):
    """pio_decode_rows_fwd/bwd against the float64 oracle of the decoder's cross_attention_layer (reference model.py:29-33,
    :229-237) with a RANDOM upstream gradient: unlike the training loss (whose decoder-query LayerNorm gradients are sums
    that cancel) every gradient here is well conditioned, so a wrong formula shows at the bf16 tolerance."""
    import perceiver
    from perceiver import functional as F
    from perceiver.model import cross_attention_layer
    from oracle import perceiver_oracle as O
    torch.manual_seed(batch + l)
    layer = cross_attention_layer(num_q_channels=e, num_kv_channels=e, num_heads=heads, dropout=0.0)
    queries = torch.nn.Parameter(torch.randn(k_queries, e) * 0.5)
    with torch.no_grad():
        for p in layer.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x_latent = torch.randn(batch, n, e)
    labels = torch.randint(0, 100, (batch, l))
    labels[torch.rand(batch, l) >= p_label] = -100
    labels[0, 0] = 5
    sel = (labels.reshape(-1) != -100).nonzero().flatten()
    dy_rows = torch.randn(sel.numel(), e)
    # oracle, float64: decode every (sample, position), take the labelled rows
    sd = {k: v.detach().double().clone().requires_grad_(True) for k, v in layer.state_dict().items()}
    q64 = queries.detach().double().clone().requires_grad_(True)
    lat64 = x_latent.double().clone().requires_grad_(True)
    full = O.cross_attention_layer(sd, "", q64[:l].unsqueeze(0).expand(batch, -1, -1), lat64, heads, None)
    ref = full.reshape(batch * l, e)[sel]
    ref.backward(dy_rows.double())
    # device
    layer = layer.to(DEV)
    qd = queries.detach().to(DEV).requires_grad_(True)
    latd = x_latent.to(DEV).requires_grad_(True)
    ca, ff = layer[0].module, layer[1].module
    a = ca.attention.attention
    with F.shadow_scope(layer):
        kv = ca.project_kv(latd)
        idx, labels_c, count = F.label_compact(labels.to(DEV), 100)
        w_packed, w_q, _, _ = a.packed()
        y = F.DecodeRowsFn.apply(qd, kv, idx, count, l, ca.q_norm.weight, ca.q_norm.bias, a.in_proj_bias, w_packed, w_q,
                                 a.out_proj.weight, a.out_proj.bias, heads, *ff._params())
    dy = torch.zeros(batch * l, e, device=DEV)
    dy[:sel.numel()] = dy_rows.to(DEV)
    y.backward(dy)
    torch.cuda.synchronize()
    assert int(count.item()) == sel.numel()
    assert_rel(y[:sel.numel()], ref, 1e-2, "decoder rows")
    assert_rel(qd.grad, q64.grad, 2e-2, "d queries")
    assert_rel(latd.grad, lat64.grad, 2e-2, "d latents")
    for k, p in layer.named_parameters():
        assert_rel(p.grad, sd[k].grad, 2e-2, f"grad {k}")
