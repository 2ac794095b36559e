"""Activations with two consumers (perceiver/functional.py, _act_pending / _act_publish): the second backward operator
adds its share of the activation's gradient into the buffer the first one returned (PIO_ACC_DX / PIO_ACC_DKV of
pio_kv_project_bwd and pio_cross_attend_bwd) instead of leaving a sum to autograd.  Checked against the same model with
the merge switched off (autograd's own summation) and, at the C ABI, against "overwrite, then add"."""
import os
from ctypes import byref

import pytest
import torch

from gpu_util import assert_rel

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(autouse=True)
def bf16_mode():
    import perceiver
    perceiver.set_compute_mode("bf16")
    yield
    perceiver.set_compute_mode("fp32")
    os.environ.pop("PIO_NO_GRAD_MERGE", None)


def _mlm(vocab, seq, latent, layers, sa, heads):
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    c = latent[1]
    enc = PerceiverEncoder(TextInputAdapter(vocab, seq, c), latent, layers, heads[0], heads[1], sa)
    dec = PerceiverDecoder(TextOutputAdapter(vocab, seq, c), latent, heads[2])
    return PerceiverMLM(enc, dec, TextMasking(vocab, 1, 2, 3))


@pytest.mark.parametrize("cfg", ["cluster", "small"])
def test_merged_activation_gradients_equal_autograds_sum(cfg):
    """3 encoder layers (layers 2 and 3 share weights AND the projected kv, reference model.py:186-191): the embedded input
    feeds two key/value projections, the shared kv two cross-attentions, and — where the out-projection rides on the
    block ("cluster" shapes) — each latent array both its cross-attention and the block behind it."""
    from perceiver import functional as F
    if cfg == "cluster":
        vocab, seq, latent, layers, sa, heads, batch = 1003, 160, (64, 256), 3, 2, (4, 4, 4), 6
    else:
        vocab, seq, latent, layers, sa, heads, batch = 503, 96, (32, 64), 3, 1, (2, 2, 2), 4
    torch.manual_seed(5)
    model = _mlm(vocab, seq, latent, layers, sa, heads).to(DEV)
    g = torch.Generator().manual_seed(1)
    ids = torch.randint(3, vocab, (batch, seq), generator=g).to(DEV)
    pad = torch.zeros(batch, seq, dtype=torch.bool, device=DEV)
    pad[1:, seq - 7:] = True
    labels = torch.full_like(ids, -100)
    sel = (torch.rand(ids.shape, device=DEV) < 0.2) & ~pad
    labels[sel] = ids[sel]

    def run(merge):
        if merge:
            os.environ.pop("PIO_NO_GRAD_MERGE", None)
        else:
            os.environ["PIO_NO_GRAD_MERGE"] = "1"
        F._ACT_STATS["merged"] = 0
        model.zero_grad(set_to_none=True)
        loss, _ = model(ids.clone(), pad, masking=False, return_loss=True, labels=labels)
        loss.backward()
        torch.cuda.synchronize()
        assert not F._PASS.acts  # the end-of-pass callback released every buffer
        return loss.detach().clone(), {k: p.grad.detach().clone() for k, p in model.named_parameters()}, F._ACT_STATS["merged"]

    loss_m, g_m, hits = run(True)
    loss_u, g_u, none = run(False)
    assert none == 0
    # x_kv (2 projections) and the shared kv (2 cross-attentions) always merge; the latent arrays when the block takes the
    # cross-attention's out-projection (3 layers)
    assert hits >= 2, hits
    if cfg == "cluster":
        assert hits == 5, hits
    assert torch.equal(loss_m, loss_u)
    for k in g_u:
        # fp32 activations: the same two terms; kv (bf16): one rounding instead of two.  Two runs of the SAME path already
        # differ by up to ~6e-3 on the small latent gradient (fp32 atomics feeding bf16 roundings), hence the bound
        assert_rel(g_m[k], g_u[k], 1.5e-2, f"grad {k}")


def test_leaf_activations_are_left_to_autograd():
    """A leaf has no grad_fn to identify it by: nothing is published for it, autograd sums."""
    from perceiver.model import PerceiverEncoder, cross_attention_layer, self_attention_block
    from perceiver.utils import Sequential
    from perceiver import functional as F
    torch.manual_seed(3)
    layer = Sequential(cross_attention_layer(256, 256, 4, 0.0), self_attention_block(2, 256, 4, 0.0)).to(DEV)
    a = torch.randn(4, 64, 256, device=DEV, requires_grad=True)
    b = torch.randn(4, 96, 256, device=DEV, requires_grad=True)
    F._ACT_STATS["merged"] = 0
    y = PerceiverEncoder._apply_layer(layer, a, b, None, None)
    y.backward(torch.randn_like(y))
    torch.cuda.synchronize()
    assert F._ACT_STATS["merged"] == 0
    assert a.grad is not None and torch.isfinite(a.grad).all()


def test_kv_project_bwd_adds_into_dx():
    """pio_kv_project_bwd(accumulate | PIO_ACC_DX): dx_kv = what it held + this projection's input gradient."""
    from perceiver import _native as N
    from perceiver import functional as F
    L = N.lib()
    torch.manual_seed(2)
    rows, ckv, e = 700, 256, 256
    x = torch.randn(rows, ckv, device=DEV)
    ln_w, ln_b = torch.randn(ckv, device=DEV), torch.randn(ckv, device=DEV)
    w = (torch.randn(2 * e, ckv, device=DEV) * 0.05).to(torch.bfloat16).contiguous()
    bias = torch.randn(2 * e, device=DEV)
    ws_ = N.attn_struct(wk=w.data_ptr(), wv=w.data_ptr() + e * ckv * 2, bk=bias.data_ptr(), bv=bias.data_ptr() + e * 4,
                        kv_ln_w=ln_w, kv_ln_b=ln_b)
    saved = torch.empty(L.pio_kv_project_saved_bytes(rows, ckv, e, N.MODE_BF16), dtype=torch.uint8, device=DEV)
    ws = torch.empty(L.pio_kv_project_ws_bytes(rows, ckv, e, N.MODE_BF16), dtype=torch.uint8, device=DEV)
    kv = torch.empty(rows, 2 * e, dtype=torch.bfloat16, device=DEV)
    st = N.stream_ptr(torch.device(DEV))
    N.check(L.pio_kv_project_fwd(x.data_ptr(), kv.data_ptr(), byref(ws_), saved.data_ptr(), rows, ckv, e, N.MODE_BF16,
                                 ws.data_ptr(), ws.numel(), st), "fwd")
    dkv = torch.randn(rows, 2 * e, device=DEV).to(torch.bfloat16)

    def bwd(dx, flags):
        dw = torch.zeros(2 * e, ckv, device=DEV)
        db = torch.zeros(2 * e, device=DEV)
        dlw, dlb = torch.zeros(ckv, device=DEV), torch.zeros(ckv, device=DEV)
        g = N.attn_struct(wk=dw.data_ptr(), wv=dw.data_ptr() + e * ckv * 4, bk=db.data_ptr(), bv=db.data_ptr() + e * 4,
                          kv_ln_w=dlw, kv_ln_b=dlb)
        rc = L.pio_kv_project_bwd(dkv.data_ptr(), x.data_ptr(), dx.data_ptr(), byref(ws_), saved.data_ptr(), byref(g), flags,
                                  rows, ckv, e, N.MODE_BF16, ws.data_ptr(), ws.numel(), st)
        torch.cuda.synchronize()
        return rc, dw, dlw

    plain = torch.empty_like(x)
    rc, dw0, dlw0 = bwd(plain, 2)
    assert rc == 0
    prev = torch.randn_like(x)
    acc = prev.clone()
    rc, dw1, dlw1 = bwd(acc, 2 | N.ACC_DX)
    assert rc == 0
    assert_rel(acc, prev + plain, 1e-6, "dx_kv")
    assert_rel(dw1, dw0, 1e-5, "dW")
    assert_rel(dlw1, dlw0, 1e-4, "d ln_w")
    # the flag is refused where it has no meaning
    rc, _, _ = bwd(acc, 0 | N.ACC_DX)
    assert rc != 0
    assert b"PIO_ACC_DX" in L.pio_last_error()


def test_cross_attend_bwd_adds_into_dkv():
    """pio_cross_attend_bwd(accumulate | PIO_ACC_DKV): dkv = bf16(what it held + this layer's key/value gradient)."""
    from perceiver import _native as N
    L = N.lib()
    torch.manual_seed(4)
    b, n, m, e, heads = 3, 64, 200, 256, 4
    assert L.pio_cross_attend_merges_dkv(b, n, m, e, heads) == 1
    x = torch.randn(b, n, e, device=DEV)
    kv = torch.randn(b, m, 2 * e, device=DEV).to(torch.bfloat16)
    pad = torch.zeros(b, m, dtype=torch.bool, device=DEV)
    pad[1, m - 9:] = True
    ln_w, ln_b = torch.randn(e, device=DEV), torch.randn(e, device=DEV)
    wq = (torch.randn(e, e, device=DEV) * 0.05).to(torch.bfloat16)
    wo = (torch.randn(e, e, device=DEV) * 0.05).to(torch.bfloat16)
    bq, bo = torch.randn(e, device=DEV), torch.randn(e, device=DEV)
    w = N.attn_struct(ln_w=ln_w, ln_b=ln_b, wq=wq, bq=bq, wo=wo, bo=bo)
    saved = torch.empty(L.pio_cross_attend_saved_bytes(b, n, m, e, heads, N.MODE_BF16), dtype=torch.uint8, device=DEV)
    ws = torch.empty(L.pio_cross_attend_ws_bytes(b, n, m, e, heads, N.MODE_BF16), dtype=torch.uint8, device=DEV)
    y = torch.empty_like(x)
    st = N.stream_ptr(torch.device(DEV))
    N.check(L.pio_cross_attend_fwd(x.data_ptr(), kv.data_ptr(), pad.data_ptr(), y.data_ptr(), byref(w), saved.data_ptr(), b, n, m, e,
                                   heads, 1, None, N.MODE_BF16, ws.data_ptr(), ws.numel(), st), "fwd")
    dy = torch.randn_like(x)

    def bwd(dkv, flags, residual=1):
        z = lambda *s: torch.zeros(*s, device=DEV)
        keep = [z(e), z(e), z(e, e), z(e), z(e, e), z(e)]
        g = N.attn_struct(ln_w=keep[0], ln_b=keep[1], wq=keep[2], bq=keep[3], wo=keep[4], bo=keep[5])
        dx = torch.empty_like(x)
        rc = L.pio_cross_attend_bwd(dy.data_ptr(), x.data_ptr(), kv.data_ptr(), pad.data_ptr(), dx.data_ptr(), dkv.data_ptr(), byref(w),
                                    saved.data_ptr(), byref(g), flags, b, n, m, e, heads, residual, None, N.MODE_BF16, ws.data_ptr(),
                                    ws.numel(), st)
        torch.cuda.synchronize()
        return rc, dx

    plain = torch.empty_like(kv)
    rc, dx0 = bwd(plain, 2)
    assert rc == 0
    prev = torch.randn(b, m, 2 * e, device=DEV).to(torch.bfloat16)
    acc = prev.clone()
    rc, dx1 = bwd(acc, 2 | N.ACC_DKV)
    assert rc == 0
    ref = prev.float() + plain.float()
    err = (acc.float() - ref).abs()
    # `plain` is itself a rounded copy of what the kernel added: half a bf16 ulp of it plus half an ulp of the sum
    assert (err <= (ref.abs() + plain.float().abs()) * 2 ** -8 + 1e-6).all(), err.max().item()
    assert_rel(dx1, dx0, 1e-6, "dx_q")
    # PIO_ACC_DX only where dx_q carries no residual term of its own
    rc, _ = bwd(acc, 2 | N.ACC_DX)
    assert rc != 0
    assert b"PIO_ACC_DX" in L.pio_last_error()
