"""pio_logits_bf16 (persistent vocabulary-logits product of the fused loss head, reference adapter.py:101-109) against
float64 torch on the same bf16 operands: full and partial tiles, a vocabulary that is no multiple of 128, device-side
row counts (none live, a few, all), zero padding up to the end of the last live row tile, untouched rows beyond."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run(m, n, k, count, off, bias=True):
    from perceiver import _native as N
    L = N.lib()
    torch.manual_seed(m + n + k)
    ldc = (n + 7) // 8 * 8
    a = torch.randn(m, k, device=DEV).bfloat16()
    w = torch.randn(n, k, device=DEV).bfloat16()
    b = torch.randn(n, device=DEV) if bias else None
    c = torch.full((m, ldc), 7.0, device=DEV, dtype=torch.bfloat16)
    cnt = None if count is None else torch.tensor([count], device=DEV, dtype=torch.int32)
    assert L.pio_logits_bf16_supported(m, n, k, k, k, ldc) == 1
    N.check(L.pio_logits_bf16(a.data_ptr(), k, w.data_ptr(), k, 0 if b is None else b.data_ptr(), c.data_ptr(), ldc, m, n, k,
                              0 if cnt is None else cnt.data_ptr(), off, N.stream_ptr(torch.device(DEV))), "pio_logits_bf16")
    torch.cuda.synchronize()
    ref = a.double() @ w.double().t() + (0 if b is None else b.double())
    live = m if count is None else max(0, min(m, count - off))
    return c, ref, live, n


@pytest.mark.parametrize("m,n,k,count,off", [(4352, 10003, 256, None, 0), (4352, 10003, 256, 4915, 4352), (4352, 10003, 256, 300, 4352),
                                             (1000, 1000, 128, None, 0), (130, 257, 64, 77, 0), (256, 128, 192, 256, 0)])
def test_logits_match_float64(m, n, k, count, off):
    c, ref, live, n = _run(m, n, k, count, off)
    if live:
        got = c[:live, :n].double()
        assert (got - ref[:live]).abs().max().item() < 6e-3 * ref[:live].abs().max().item()  # one bf16 rounding of the result
    pad_end = min(c.shape[0], (live + 127) // 128 * 128)
    if pad_end > live:
        assert (c[live:pad_end, :n] == 0).all()  # operand padding for the product over the row axis
    if pad_end < c.shape[0]:
        assert (c[pad_end:] == 7.0).all()  # rows beyond the live tiles are not touched


def test_logits_nothing_live_touches_nothing():
    c, ref, live, n = _run(4352, 10003, 256, 4000, 4352)
    assert live == 0 and (c == 7.0).all()


def test_logits_without_bias_and_unsupported_shapes():
    from perceiver import _native as N
    c, ref, live, n = _run(256, 384, 256, None, 0, bias=False)
    assert (c[:, :n].double() - ref).abs().max().item() < 6e-3 * ref.abs().max().item()
    L = N.lib()
    assert L.pio_logits_bf16_supported(256, 384, 320, 320, 320, 384) == 0  # k > 256
    assert L.pio_logits_bf16_supported(256, 64, 256, 256, 256, 64) == 0    # n < 128
