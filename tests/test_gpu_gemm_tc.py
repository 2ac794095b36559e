"""tcgen05 bf16 GEMM (pio_gemm_bf16) against an fp64 product of the same bf16-rounded operands —
the only error left is fp32 accumulation order, so the tolerance is tight (1e-5 of the output scale);
bf16-rounded outputs get bf16 tolerance."""
import ctypes
import math

import pytest
import torch

from gpu_util import assert_rel
from oracle import perceiver_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def N():
    from perceiver import _native
    _native.lib()
    return _native


def _pad8(x):
    return (x + 7) // 8 * 8


def _bf16_padded(t, ld):
    """(rows, cols) fp32 -> bf16 device tensor with pitch ld (garbage-free padding)"""
    out = torch.zeros(t.shape[0], ld, dtype=torch.bfloat16, device=DEV)
    out[:, :t.shape[1]] = t.to(DEV).to(torch.bfloat16)
    return out


def _run(N, a, b, m, n, k, a_k, b_k, **kw):
    L = N.lib()
    d = N.GemmBf16Desc()
    d.a, d.b = a.data_ptr(), b.data_ptr()
    d.m, d.n, d.k = m, n, k
    d.lda, d.ldb = a.shape[1], b.shape[1]
    d.a_kmajor, d.b_kmajor = int(a_k), int(b_k)
    c32 = kw.get("c32")
    if c32 is None and not kw.get("no_c32"):
        c32 = torch.full((m, n), float("nan"), device=DEV)
    c16 = kw.get("c16")
    d.c32, d.ldc32 = N.ptr(c32), n
    d.c16, d.ldc16 = N.ptr(c16), (c16.shape[1] if c16 is not None else 0)
    d.bias = N.ptr(kw.get("bias"))
    d.residual, d.ldr = N.ptr(kw.get("residual")), n
    aux = kw.get("aux16")
    d.aux16, d.ldaux = N.ptr(aux), (aux.shape[1] if aux is not None else 0)
    d.act = kw.get("act", 0)
    d.accumulate = kw.get("accumulate", 0)
    ws = torch.empty(L.pio_gemm_bf16_ws_bytes(m, n, k), dtype=torch.uint8, device=DEV)
    N.check(L.pio_gemm_bf16(ctypes.byref(d), ws.data_ptr(), ws.numel(), N.stream_ptr()), "pio_gemm_bf16")
    torch.cuda.synchronize()
    return c32


SHAPES = [(128, 128, 64), (128, 64, 64), (256, 256, 256), (4096, 256, 256), (100, 10, 131), (1, 7, 5), (300, 768, 256),
          (129, 130, 17), (1000, 10003, 64), (32768, 512, 256)]


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_tc_gemm_kmajor_kmajor(N, m, n, k):
    """forward layout: X (m,k) W (n,k)"""
    torch.manual_seed(0)
    x, w = torch.randn(m, k), torch.randn(n, k) / math.sqrt(k)
    xb, wb = _bf16_padded(x, _pad8(k)), _bf16_padded(w, _pad8(k))
    ref = xb[:, :k].double().cpu() @ wb[:, :k].double().cpu().t()
    c = _run(N, xb, wb, m, n, k, True, True)
    assert_rel(c, ref, 1e-5, "NT")


@pytest.mark.parametrize("m,n,k", SHAPES)
def test_tc_gemm_kmajor_mnmajor(N, m, n, k):
    """input-gradient layout: dY (m,k) K-major, W stored (k, n) i.e. B MN-major"""
    torch.manual_seed(1)
    dy, w = torch.randn(m, k), torch.randn(k, n) / math.sqrt(k)
    dyb, wb = _bf16_padded(dy, _pad8(k)), _bf16_padded(w, _pad8(n))
    ref = dyb[:, :k].double().cpu() @ wb[:, :n].double().cpu()
    c = _run(N, dyb, wb, m, n, k, True, False)
    assert_rel(c, ref, 1e-5, "NN")


@pytest.mark.parametrize("m,n,k", [(256, 256, 4096), (768, 256, 8192), (10, 131, 1000), (5, 3, 77), (512, 256, 32768),
                                   (130, 129, 300), (10003, 256, 2048)])
def test_tc_gemm_mnmajor_mnmajor_splitk(N, m, n, k):
    """weight-gradient layout: dY stored (k, m), X stored (k, n): both MN-major; split-K"""
    torch.manual_seed(2)
    dy, x = torch.randn(k, m), torch.randn(k, n)
    dyb, xb = _bf16_padded(dy, _pad8(m)), _bf16_padded(x, _pad8(n))
    ref = dyb[:, :m].double().cpu().t() @ xb[:, :n].double().cpu()
    c = _run(N, dyb, xb, m, n, k, False, False)
    assert_rel(c, ref, 1e-5, "TN")
    c2 = _run(N, dyb, xb, m, n, k, False, False)
    assert torch.equal(c, c2)  # deterministic split-K
    base = torch.ones(m, n, device=DEV)
    c3 = _run(N, dyb, xb, m, n, k, False, False, c32=base, accumulate=1)
    assert_rel(c3, ref + 1, 1e-5, "TN accumulate")


def test_tc_gemm_mnmajor_kmajor(N):
    torch.manual_seed(3)
    m, n, k = 200, 96, 136
    a, b = torch.randn(k, m), torch.randn(n, k)
    ab, bb = _bf16_padded(a, _pad8(m)), _bf16_padded(b, _pad8(k))
    ref = ab[:, :m].double().cpu().t() @ bb[:, :k].double().cpu().t()
    assert_rel(_run(N, ab, bb, m, n, k, False, True), ref, 1e-5, "TT")


@pytest.mark.parametrize("m,n,k", [(4096, 256, 256), (100, 10, 131), (333, 200, 72)])
def test_tc_gemm_epilogues(N, m, n, k):
    torch.manual_seed(4)
    x, w = torch.randn(m, k), torch.randn(n, k) / math.sqrt(k)
    bias, res = torch.randn(n), torch.randn(m, n)
    xb, wb = _bf16_padded(x, _pad8(k)), _bf16_padded(w, _pad8(k))
    ref = xb[:, :k].double().cpu() @ wb[:, :k].double().cpu().t()
    c = _run(N, xb, wb, m, n, k, True, True, bias=bias.to(DEV), residual=res.to(DEV))
    assert_rel(c, ref + bias.double() + res.double(), 1e-5, "bias+res")
    # GELU with the derivative saved as bf16 (one erf for both) and bf16 output
    ldn = _pad8(n)
    aux = torch.zeros(m, ldn, dtype=torch.bfloat16, device=DEV)
    c16 = torch.zeros(m, ldn, dtype=torch.bfloat16, device=DEV)
    c = _run(N, xb, wb, m, n, k, True, True, bias=bias.to(DEV), aux16=aux, act=1, c16=c16)
    pre = (ref + bias.double()).requires_grad_(True)
    O.gelu_erf(pre).sum().backward()
    assert_rel(aux[:, :n].float(), pre.grad, 8e-3, "gelu' bf16")
    pre = pre.detach()
    assert_rel(c, O.gelu_erf(pre), 1e-5, "gelu fp32")
    assert_rel(c16[:, :n].float(), O.gelu_erf(pre), 8e-3, "gelu bf16")
    assert (c16[:, n:] == 0).all()
    # act=2 multiplies by the saved derivative
    c = _run(N, xb, wb, m, n, k, True, True, aux16=aux, act=2)
    assert_rel(c, ref * aux[:, :n].double().cpu(), 2e-5, "dgelu")
    # bf16-only output
    c16.zero_()
    _run(N, xb, wb, m, n, k, True, True, c16=c16, no_c32=True)
    assert_rel(c16[:, :n].float(), ref, 8e-3, "bf16 only")


def test_cast_bf16_with_padding(N):
    x = torch.randn(37, 131, device=DEV)
    out = torch.full((37, 136), 7.0, dtype=torch.bfloat16, device=DEV)
    N.check(N.lib().pio_cast_bf16(x.data_ptr(), out.data_ptr(), 37, 131, 131, 136, N.stream_ptr()))
    assert torch.equal(out[:, :131], x.to(torch.bfloat16)) and (out[:, 131:] == 0).all()


def test_tc_gemm_rejects_misaligned_operands(N):
    a = torch.zeros(16, 20, dtype=torch.bfloat16, device=DEV)  # pitch 20 is not a multiple of 8
    b = torch.zeros(16, 24, dtype=torch.bfloat16, device=DEV)
    with pytest.raises(RuntimeError, match="multiple of 8"):
        _run(N, a, b, 16, 16, 20, True, True)
