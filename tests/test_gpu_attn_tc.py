"""tcgen05 flash attention forward (pio_attn_fwd, PIO_MODE_BF16, dh = 32 / 64 / 128) against an fp64 softmax
attention over the same bf16-rounded q/k/v.  P is rounded to bf16 before the PV product, so the
output tolerance is the bf16 one (1e-2 of the output scale; measured ~2e-3); lse is fp32-accurate."""
import ctypes

import pytest
import torch

from gpu_util import BF16_TOL, assert_rel

pytestmark = pytest.mark.gpu
DEV = "cuda"


@pytest.fixture(scope="module")
def N():
    from perceiver import _native
    _native.lib()
    return _native


def _ref(q, k, v, pad, heads):
    b, n, e = q.shape
    m = k.shape[1]
    dh = e // heads
    qh = q.view(b, n, heads, dh).transpose(1, 2)
    kh = k.view(b, m, heads, dh).transpose(1, 2)
    vh = v.view(b, m, heads, dh).transpose(1, 2)
    s = (qh @ kh.transpose(-1, -2)) * dh ** -0.5
    if pad is not None:
        s = s.masked_fill(pad[:, None, None, :], float("-inf"))
    return (torch.softmax(s, -1) @ vh).transpose(1, 2).reshape(b, n, e), torch.logsumexp(s, -1)


def _run(N, q16, k16, v16, q_rs, k_rs, pad, b, heads, n, m, dh, mode=1):
    e = heads * dh
    o = torch.full((b, n, e), float("nan"), device=DEV)
    lse = torch.full((b, heads, n), float("nan"), device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = q16.data_ptr(), k16.data_ptr(), v16.data_ptr(), N.ptr(pad), o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * q_rs, q_rs, m * k_rs, k_rs, m * k_rs, k_rs, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    N.check(N.lib().pio_attn_fwd(ctypes.byref(d), mode, N.stream_ptr()), "pio_attn_fwd")
    torch.cuda.synchronize()
    return o, lse


@pytest.mark.parametrize("dh", [32, 64, 128])
@pytest.mark.parametrize("b,n,m,heads,padded", [
    (2, 128, 64, 1, False), (2, 128, 128, 2, False), (3, 64, 512, 4, True), (2, 256, 1024, 8, False), (2, 70, 200, 2, True),
    (1, 1, 64, 1, False), (2, 512, 64, 4, False), (1, 300, 4100, 2, True), (2, 64, 64, 4, False)])
def test_flash_tc_forward_cross_layout(N, b, n, m, heads, padded, dh):
    """separate q (B,N,E) and packed kv (B,M,2E) — the cross-attention layout; head sizes 64 and 128"""
    torch.manual_seed(0)
    e = heads * dh
    q = torch.randn(b, n, e)
    kv = torch.randn(b, m, 2 * e)
    pad = None
    if padded:
        lens = torch.randint(1, m + 1, (b,))
        lens[0] = m
        pad = torch.arange(m)[None, :] >= lens[:, None]
    q16, kv16 = q.to(DEV).to(torch.bfloat16), kv.to(DEV).to(torch.bfloat16)
    o_ref, lse_ref = _ref(q16.double().cpu(), kv16[..., :e].double().cpu(), kv16[..., e:].double().cpu(), pad, heads)
    o, lse = _run(N, q16, kv16, kv16[..., e:], e, 2 * e, pad.to(DEV) if pad is not None else None, b, heads, n, m, dh)
    assert_rel(o, o_ref, BF16_TOL, "o")
    assert_rel(lse, lse_ref, 1e-4, "lse")


@pytest.mark.parametrize("b,n,heads", [(4, 64, 4), (2, 256, 2), (3, 100, 1)])
def test_flash_tc_forward_packed_qkv_layout(N, b, n, heads):
    """packed (B,N,3C) qkv — the self-attention layout"""
    torch.manual_seed(1)
    dh = 64
    c = heads * dh
    qkv = torch.randn(b, n, 3 * c)
    x16 = qkv.to(DEV).to(torch.bfloat16)
    xr = x16.double().cpu()
    o_ref, lse_ref = _ref(xr[..., :c], xr[..., c:2 * c], xr[..., 2 * c:], None, heads)
    o, lse = _run(N, x16, x16[..., c:], x16[..., 2 * c:], 3 * c, 3 * c, None, b, heads, n, n, dh)
    assert_rel(o, o_ref, BF16_TOL, "o")
    assert_rel(lse, lse_ref, 1e-4, "lse")


def test_flash_tc_all_padded_row_is_nan(N):
    b, n, m, heads, dh = 2, 8, 100, 1, 64
    q16 = torch.randn(b, n, 64, device=DEV).to(torch.bfloat16)
    kv16 = torch.randn(b, m, 128, device=DEV).to(torch.bfloat16)
    pad = torch.zeros(b, m, dtype=torch.bool, device=DEV)
    pad[1] = True
    o, _ = _run(N, q16, kv16, kv16[..., 64:], 64, 128, pad, b, heads, n, m, dh)
    assert torch.isfinite(o[0]).all() and torch.isnan(o[1]).all()


def test_bf16_inputs_other_head_sizes_use_the_fp32_arithmetic_kernel(N):
    torch.manual_seed(2)
    b, n, m, heads, dh = 2, 20, 50, 2, 16  # (32, 64 and 128 run on the tensor-core kernels)
    e = heads * dh
    q16 = torch.randn(b, n, e, device=DEV).to(torch.bfloat16)
    kv16 = torch.randn(b, m, 2 * e, device=DEV).to(torch.bfloat16)
    o_ref, lse_ref = _ref(q16.double().cpu(), kv16[..., :e].double().cpu(), kv16[..., e:].double().cpu(), None, heads)
    o, lse = _run(N, q16, kv16, kv16[..., e:], e, 2 * e, None, b, heads, n, m, dh)
    assert_rel(o, o_ref, 1e-5, "o")
    assert_rel(lse, lse_ref, 1e-5, "lse")


def _ref_grads(q, k, v, do, pad, heads):
    q, k, v = (t.clone().requires_grad_(True) for t in (q, k, v))
    o, lse = _ref(q, k, v, pad, heads)
    (o * do).sum().backward()
    return o.detach(), lse.detach(), q.grad, k.grad, v.grad


@pytest.mark.parametrize("dh", [32, 64, 128])
@pytest.mark.parametrize("b,n,m,heads,padded", [
    (2, 128, 64, 1, False), (2, 64, 512, 4, True), (2, 256, 1024, 2, False), (2, 70, 200, 2, True), (1, 1, 64, 1, False),
    (2, 512, 64, 4, False), (1, 300, 1100, 2, True), (3, 64, 64, 4, False), (2, 256, 256, 4, False), (2, 100, 1100, 2, True),
    (1, 128, 300, 3, False)])
@pytest.mark.parametrize("one_tile_kernel", [True, False])
def test_flash_tc_backward(N, b, n, m, heads, padded, dh, one_tile_kernel, monkeypatch):
    """dq, dk, dv of the tcgen05 backward against fp64 autograd over the same bf16-rounded q, k, v, dO.
    The kernels round P, dS and the outputs to bf16, hence the bf16 tolerance.  Head sizes 64 and 128.
    n <= 128 against more than 128 keys at head size 64 runs the (sample, head) kernel of pio_xattn_tc.cu unless switched off; everything else the
    two general kernels."""
    one_tile = n <= 128 and m > 128 and dh == 64  # (keys that fit one tile stay with the general kernels)
    if not one_tile_kernel:
        if not one_tile:
            pytest.skip("same path as one_tile_kernel=True")
        monkeypatch.setenv("PIO_NO_XATTN_TC", "1")
    torch.manual_seed(3)
    e = heads * dh
    q, kv, do = torch.randn(b, n, e), torch.randn(b, m, 2 * e), torch.randn(b, n, e)
    pad = None
    if padded:
        lens = torch.randint(1, m + 1, (b,))
        lens[0] = m
        pad = torch.arange(m)[None, :] >= lens[:, None]
    q16, kv16, do16 = (t.to(DEV).to(torch.bfloat16) for t in (q, kv, do))
    o_ref, lse_ref, dq_ref, dk_ref, dv_ref = _ref_grads(q16.double().cpu(), kv16[..., :e].double().cpu(),
                                                        kv16[..., e:].double().cpu(), do16.double().cpu(), pad, heads)
    padd = pad.to(DEV) if pad is not None else None
    # forward through the tensor-core kernel (its fp32 o and lse are what the backward consumes)
    L = N.lib()
    o = torch.empty(b, n, e, device=DEV)
    lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = q16.data_ptr(), kv16.data_ptr(), kv16.data_ptr() + 2 * e, N.ptr(padd), o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * 2 * e, 2 * e, m * 2 * e, 2 * e, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    N.check(L.pio_attn_fwd(ctypes.byref(d), 1, N.stream_ptr()), "pio_attn_fwd")
    dq16 = torch.full((b, n, e), float("nan"), device=DEV, dtype=torch.bfloat16)
    dkv16 = torch.full((b, m, 2 * e), float("nan"), device=DEV, dtype=torch.bfloat16)
    delta = torch.empty(b, heads, n, device=DEV)
    N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do16.data_ptr(), dq16.data_ptr(), dkv16.data_ptr(), dkv16.data_ptr() + 2 * e,
                                delta.data_ptr(), N.stream_ptr()), "pio_attn_bwd_bf16")
    torch.cuda.synchronize()
    if not (one_tile and one_tile_kernel):  # (the one-tile kernel forms delta in registers)
        assert_rel(delta, (o_ref * do16.double().cpu()).view(b, n, heads, dh).sum(-1).transpose(1, 2), 2e-2, "delta")
    assert_rel(dq16.float(), dq_ref, 2e-2, "dq")
    assert_rel(dkv16[..., :e].float(), dk_ref, 2e-2, "dk")
    assert_rel(dkv16[..., e:].float(), dv_ref, 2e-2, "dv")


def test_flash_tc_backward_packed_qkv(N):
    torch.manual_seed(4)
    b, n, heads, dh = 3, 100, 2, 64
    c = heads * dh
    x16 = torch.randn(b, n, 3 * c, device=DEV).to(torch.bfloat16)
    do16 = torch.randn(b, n, c, device=DEV).to(torch.bfloat16)
    xr = x16.double().cpu()
    _, _, dq_ref, dk_ref, dv_ref = _ref_grads(xr[..., :c], xr[..., c:2 * c], xr[..., 2 * c:], do16.double().cpu(), None, heads)
    L = N.lib()
    o = torch.empty(b, n, c, device=DEV)
    lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = x16.data_ptr(), x16.data_ptr() + 2 * c, x16.data_ptr() + 4 * c, None, o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * 3 * c, 3 * c, n * 3 * c, 3 * c, n * 3 * c, 3 * c, n * c, c
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, n, dh, dh ** -0.5
    N.check(L.pio_attn_fwd(ctypes.byref(d), 1, N.stream_ptr()))
    dx16 = torch.full((b, n, 3 * c), float("nan"), device=DEV, dtype=torch.bfloat16)
    delta = torch.empty(b, heads, n, device=DEV)
    N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do16.data_ptr(), dx16.data_ptr(), dx16.data_ptr() + 2 * c, dx16.data_ptr() + 4 * c,
                                delta.data_ptr(), N.stream_ptr()))
    torch.cuda.synchronize()
    assert_rel(dx16[..., :c].float(), dq_ref, 2e-2, "dq")
    assert_rel(dx16[..., c:2 * c].float(), dk_ref, 2e-2, "dk")
    assert_rel(dx16[..., 2 * c:].float(), dv_ref, 2e-2, "dv")


@pytest.mark.parametrize("b,n,m,heads,dh", [(1, 256, 16384, 8, 64), (1, 1, 8192, 2, 64), (1, 128, 16384, 2, 128)])
def test_flash_tc_forward_and_backward_at_long_key_axes(N, b, n, m, heads, dh):
    """The north-star shape class (M >= 8192; BASELINE configs[3]: N 256, M 16384, 8 heads of 64) and the N = 1 decoder
    query against fp64: the forward's lazy rescaling (the reference maximum may lag by 2^8) and the backward's tile-by-tile
    accumulation in TMEM over 256 key tiles."""
    torch.manual_seed(9)
    e = heads * dh
    q16 = torch.randn(b, n, e, device=DEV).to(torch.bfloat16)
    kv16 = (torch.randn(b, m, 2 * e, device=DEV) * 1.5).to(torch.bfloat16)  # wider scores: the running maximum really moves
    do16 = torch.randn(b, n, e, device=DEV).to(torch.bfloat16)
    o_ref, lse_ref, dq_ref, dk_ref, dv_ref = _ref_grads(q16.double().cpu(), kv16[..., :e].double().cpu(), kv16[..., e:].double().cpu(),
                                                        do16.double().cpu(), None, heads)
    L = N.lib()
    o = torch.empty(b, n, e, device=DEV)
    lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = q16.data_ptr(), kv16.data_ptr(), kv16.data_ptr() + 2 * e, None, o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * 2 * e, 2 * e, m * 2 * e, 2 * e, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    N.check(L.pio_attn_fwd(ctypes.byref(d), 1, N.stream_ptr()), "pio_attn_fwd")
    dq16 = torch.full((b, n, e), float("nan"), device=DEV, dtype=torch.bfloat16)
    dkv16 = torch.full((b, m, 2 * e), float("nan"), device=DEV, dtype=torch.bfloat16)
    delta = torch.empty(b, heads, n, device=DEV)
    N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do16.data_ptr(), dq16.data_ptr(), dkv16.data_ptr(), dkv16.data_ptr() + 2 * e,
                                delta.data_ptr(), N.stream_ptr()), "pio_attn_bwd_bf16")
    torch.cuda.synchronize()
    assert_rel(o, o_ref, BF16_TOL, "o")
    assert_rel(lse, lse_ref, 1e-4, "lse")
    assert_rel(dq16.float(), dq_ref, 2e-2, "dq")
    assert_rel(dkv16[..., :e].float(), dk_ref, 2e-2, "dk")
    assert_rel(dkv16[..., e:].float(), dv_ref, 2e-2, "dv")
