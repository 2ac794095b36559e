"""pio_dw_batched (one launch for a table of weight-gradient products, bias gradients riding on it) and
pio_ln_param_grads (LayerNorm dgamma / dbeta for a table of LayerNorms) against float64 torch on the same bf16 inputs.
Both serve the fused backward chain of the latent block (reference: autograd through self_attention_layer,
model.py:20-26, :36-44, :102-116)."""
import ctypes

import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


class DwItem(ctypes.Structure):
    _fields_ = [("dy16", ctypes.c_void_p), ("lddy", ctypes.c_int64), ("m", ctypes.c_int), ("x16", ctypes.c_void_p),
                ("ldx", ctypes.c_int64), ("n", ctypes.c_int), ("dw", ctypes.c_void_p), ("db", ctypes.c_void_p)]


class LnItem(ctypes.Structure):
    _fields_ = [("dn16", ctypes.c_void_p), ("x", ctypes.c_void_p), ("stats", ctypes.c_void_p), ("dgamma", ctypes.c_void_p),
                ("dbeta", ctypes.c_void_p)]


@pytest.mark.parametrize("rows,ms", [(4096, [256, 256, 768]), (200, [256]), (1000, [768, 256, 256, 256, 768, 8]), (64, [128] * 7)])
def test_dw_batched_matches_float64(rows, ms):
    from perceiver import _native as N
    torch.manual_seed(rows + len(ms))
    L = N.lib()
    items = (DwItem * len(ms))()
    keep, refs = [], []
    for i, m in enumerate(ms):
        dy = torch.randn(rows, m, device=DEV).bfloat16()
        x = torch.randn(rows, 256, device=DEV).bfloat16()
        dw = torch.randn(m, 256, device=DEV)  # running gradients: the kernel accumulates
        db = torch.randn(m, device=DEV) if i % 2 == 0 else None
        refs.append((dw.double() + dy.double().t() @ x.double(), None if db is None else db.double() + dy.double().sum(0)))
        items[i] = DwItem(dy.data_ptr(), m, m, x.data_ptr(), 256, 256, dw.data_ptr(), 0 if db is None else db.data_ptr())
        keep.append((dy, x, dw, db))
    N.check(L.pio_dw_batched(items, len(ms), rows, N.stream_ptr(torch.device(DEV))), "pio_dw_batched")
    torch.cuda.synchronize()
    for (dy, x, dw, db), (rw, rb) in zip(keep, refs):
        scale = rw.abs().max().item()
        assert (dw.double() - rw).abs().max().item() < 2e-5 * scale + 1e-4, "dw"
        if db is not None:
            assert (db.double() - rb).abs().max().item() < 2e-5 * rb.abs().max().item() + 1e-4, "db"


def test_dw_batched_rejects_unsupported_items():
    from perceiver import _native as N
    L = N.lib()
    dy = torch.zeros(64, 64, device=DEV, dtype=torch.bfloat16)
    x = torch.zeros(64, 128, device=DEV, dtype=torch.bfloat16)
    dw = torch.zeros(64, 128, device=DEV)
    items = (DwItem * 1)(DwItem(dy.data_ptr(), 64, 64, x.data_ptr(), 128, 128, dw.data_ptr(), 0))
    assert L.pio_dw_batched(items, 1, 64, N.stream_ptr(torch.device(DEV))) != 0  # n must be 256
    assert L.pio_dw_batched(items, 0, 64, N.stream_ptr(torch.device(DEV))) != 0


@pytest.mark.parametrize("rows,c,count", [(4096, 256, 12), (130, 256, 1), (777, 64, 3)])
def test_ln_param_grads_match_float64(rows, c, count):
    from perceiver import _native as N
    torch.manual_seed(rows + c)
    L = N.lib()
    items = (LnItem * count)()
    keep, refs = [], []
    for i in range(count):
        dn = torch.randn(rows, c, device=DEV).bfloat16()
        x = torch.randn(rows, c, device=DEV) * 2 + 0.5
        mean = x.mean(1)
        rstd = (x.var(1, unbiased=False) + 1e-5).rsqrt()
        stats = torch.stack([mean, rstd], 1).contiguous()
        dg = torch.randn(c, device=DEV)
        db = torch.randn(c, device=DEV)
        xhat = (x.double() - mean.double()[:, None]) * rstd.double()[:, None]
        refs.append((dg.double() + (dn.double() * xhat).sum(0), db.double() + dn.double().sum(0)))
        items[i] = LnItem(dn.data_ptr(), x.data_ptr(), stats.data_ptr(), dg.data_ptr(), db.data_ptr())
        keep.append((dn, x, stats, dg, db))
    N.check(L.pio_ln_param_grads(items, count, rows, c, N.stream_ptr(torch.device(DEV))), "pio_ln_param_grads")
    torch.cuda.synchronize()
    for (dn, x, stats, dg, db), (rg, rb) in zip(keep, refs):
        assert (dg.double() - rg).abs().max().item() < 1e-4 * rg.abs().max().item() + 1e-4
        assert (db.double() - rb).abs().max().item() < 1e-4 * rb.abs().max().item() + 1e-4
