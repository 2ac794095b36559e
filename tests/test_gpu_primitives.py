"""Parity of each CUDA entry point (through the C ABI) against the CPU oracle, on seeded inputs."""
import ctypes
import math

import pytest
import torch

from conftest import load_golden
from gpu_util import FP32_TOL, assert_rel, rel_err
from oracle import perceiver_oracle as O

pytestmark = pytest.mark.gpu

DEV = "cuda"


@pytest.fixture(scope="module")
def N():
    from perceiver import _native
    _native.lib()
    return _native


def test_image_encode_bit_exact(N):
    from perceiver import ImageInputAdapter
    for name in ("fourier_6x5_f3", "fourier_3d"):
        meta, a, _, _ = load_golden(name)
        ad = ImageInputAdapter(tuple(meta["image_shape"]), meta["bands"]).to(DEV)
        out = ad(a["x"].to(DEV))
        assert torch.equal(out.cpu(), a["out"])
    # MNIST shape (BASELINE configs[0]) against the oracle, full batch
    torch.manual_seed(0)
    x = torch.rand(128, 28, 28, 1) * 2 - 1
    ad = ImageInputAdapter((28, 28, 1), 32).to(DEV)
    ref = O.image_input_adapter({"position_encoding": ad.position_encoding.cpu()}, "", x)
    got = ad(x.to(DEV))
    assert got.shape == (128, 784, 131)
    assert torch.equal(got.cpu(), ref)
    # empty batch
    assert ad(torch.empty(0, 28, 28, 1, device=DEV)).shape == (0, 784, 131)


def test_fourier_table_on_device(N):
    from perceiver import ImageInputAdapter
    for shape, bands in (((28, 28, 1), 32), ((6, 5, 1), 3), ((3, 4, 2, 2), 2), ((64, 48, 3), 16)):
        ad = ImageInputAdapter(shape, bands)
        host = ad.position_encoding.clone()
        ad = ad.to(DEV).regenerate_on_device()
        d = len(shape) - 1
        assert torch.equal(ad.position_encoding[:, :d].cpu(), host[:, :d])       # positions exact
        assert (ad.position_encoding.cpu() - host).abs().max().item() < 5e-6      # sin/cos within ulps


def test_text_embed_fwd_bwd(N):
    from perceiver import TextInputAdapter
    torch.manual_seed(1)
    vocab, seq, c, b, l = 101, 24, 40, 5, 17
    ad = TextInputAdapter(vocab, seq, c)
    sd = {k: v.detach().clone().requires_grad_(True) for k, v in ad.state_dict().items()}
    ids = torch.randint(0, vocab, (b, l))
    ids[0, :3] = 7  # repeated ids exercise the atomic accumulation
    w = torch.randn(b, l, c)
    ref = O.text_input_adapter(sd, "", ids)
    (ref * w).sum().backward()
    ad = ad.to(DEV)
    got = ad(ids.to(DEV))
    (got * w.to(DEV)).sum().backward()
    assert_rel(got, ref, 1e-6, "embed fwd")
    assert_rel(ad.text_embedding.weight.grad, sd["text_embedding.weight"].grad, 1e-5, "demb")
    assert_rel(ad.pos_encoding.grad, sd["pos_encoding"].grad, 1e-5, "dpos")
    with pytest.raises(ValueError):
        ad(torch.zeros(2, seq + 1, dtype=torch.long, device=DEV))


@pytest.mark.parametrize("shape", [(8, 96), (64, 512), (3, 1000), (1, 1), (2, 1025)])
def test_text_masking_bit_exact_vs_oracle_same_generator(N, shape):
    from perceiver import TextMasking
    vocab = 10003
    g0 = torch.Generator().manual_seed(11)
    ids = torch.randint(3, vocab, shape, generator=g0)
    lens = torch.randint(1, shape[1] + 1, (shape[0],), generator=g0)
    pad = torch.arange(shape[1])[None, :] >= lens[:, None]
    ids[torch.rand(shape, generator=g0) < 0.03] = 1
    ids[pad] = 0
    ids, pad = ids.to(DEV), pad.to(DEV)
    ga = torch.Generator(device=DEV).manual_seed(5)
    gb = torch.Generator(device=DEV).manual_seed(5)
    x_ref, lab_ref = O.text_masking(ids.clone(), pad, vocab_size=vocab, unk_token_id=1, mask_token_id=2,
                                    num_special_tokens=3, generator=ga)
    x_in = ids.clone()
    x_got, lab_got = TextMasking(vocab, 1, 2, 3)(x_in, pad, generator=gb)
    assert x_got.data_ptr() == x_in.data_ptr()  # in place, like the reference
    assert torch.equal(x_got, x_ref)
    assert torch.equal(lab_got, lab_ref)
    assert torch.equal(ga.get_state(), gb.get_state())  # consumed the same draws
    # global-generator path too (what the reference itself uses)
    torch.manual_seed(123)
    x_ref2, lab_ref2 = O.text_masking(ids.clone(), pad, vocab_size=vocab, unk_token_id=1, mask_token_id=2,
                                      num_special_tokens=3)
    torch.manual_seed(123)
    x_got2, lab_got2 = TextMasking(vocab, 1, 2, 3)(ids.clone(), pad)
    assert torch.equal(x_got2, x_ref2) and torch.equal(lab_got2, lab_ref2)


def test_text_masking_statistics_full_size(N):
    from perceiver import TextMasking
    torch.manual_seed(0)
    ids = torch.randint(3, 10003, (64, 512), device=DEV)
    pad = torch.zeros(64, 512, dtype=torch.bool, device=DEV)
    orig = ids.clone()
    x, lab = TextMasking(10003, 1, 2, 3)(ids, pad)
    sel = lab != -100
    frac = sel.float().mean().item()
    assert 0.14 < frac < 0.16
    assert torch.equal(lab[sel], orig[sel])
    masked = (x == 2) & sel
    assert 0.78 < masked.sum().item() / sel.sum().item() < 0.82
    assert torch.equal(x[~sel], orig[~sel])


@pytest.mark.parametrize("rows,c", [(37, 16), (1000, 131), (4096, 256), (513, 512), (5, 1000)])
def test_layernorm_fwd_bwd(N, rows, c):
    torch.manual_seed(2)
    x = torch.randn(rows, c) * 2 + 0.5
    w = torch.randn(c)
    b = torch.randn(c)
    dy = torch.randn(rows, c)
    dres = torch.randn(rows, c)
    xr, wr, br = (t.double().requires_grad_(True) for t in (x, w, b))
    yr = O.layer_norm(xr, wr, br)
    (yr * dy.double()).sum().backward()
    L = N.lib()
    xd, wd, bd, dyd, dresd = (t.to(DEV) for t in (x, w, b, dy, dres))
    y = torch.empty_like(xd)
    stats = torch.empty(rows, 2, device=DEV)
    st = N.stream_ptr()
    N.check(L.pio_layernorm_fwd(xd.data_ptr(), wd.data_ptr(), bd.data_ptr(), y.data_ptr(), stats.data_ptr(), rows, c, 1e-5, st))
    assert_rel(y, yr, 1e-5, "ln fwd")
    dx = torch.empty_like(xd)
    dw = torch.empty(c, device=DEV)
    db = torch.empty(c, device=DEV)
    ws = torch.empty(L.pio_layernorm_bwd_ws_bytes(rows, c), dtype=torch.uint8, device=DEV)
    N.check(L.pio_layernorm_bwd(dyd.data_ptr(), xd.data_ptr(), stats.data_ptr(), wd.data_ptr(), dresd.data_ptr(),
                                dx.data_ptr(), dw.data_ptr(), db.data_ptr(), 0, rows, c, ws.data_ptr(), ws.numel(), st))
    assert_rel(dx, xr.grad + dres.double(), 1e-5, "ln dx")
    assert_rel(dw, wr.grad, 1e-5, "ln dw")
    assert_rel(db, br.grad, 1e-5, "ln db")
    # accumulate=1 adds on top
    N.check(L.pio_layernorm_bwd(dyd.data_ptr(), xd.data_ptr(), stats.data_ptr(), wd.data_ptr(), None,
                                None, dw.data_ptr(), db.data_ptr(), 1, rows, c, ws.data_ptr(), ws.numel(), st))
    assert_rel(dw, 2 * wr.grad, 1e-5, "ln dw acc")


def _gemm(N, a, b, m, n, k, a_k, b_k, bias=None, residual=None, aux=None, act=0, accumulate=0, c=None, rowsum=False):
    L = N.lib()
    d = N.GemmDesc()
    c = torch.empty(m, n, device=DEV) if c is None else c
    d.a, d.b, d.c = a.data_ptr(), b.data_ptr(), c.data_ptr()
    d.m, d.n, d.k = m, n, k
    d.lda, d.ldb, d.ldc = a.shape[1], b.shape[1], n
    d.a_kmajor, d.b_kmajor = int(a_k), int(b_k)
    d.bias = N.ptr(bias)
    d.residual = N.ptr(residual)
    d.ldr = n
    d.aux = N.ptr(aux)
    d.ldaux = n
    d.act = act
    d.accumulate = accumulate
    rs = torch.zeros(m, device=DEV) if rowsum else None
    d.rowsum_a = N.ptr(rs)
    ws = torch.empty(L.pio_gemm_f32_ws_bytes(m, n, k), dtype=torch.uint8, device=DEV)
    N.check(L.pio_gemm_f32(ctypes.byref(d), ws.data_ptr(), ws.numel(), N.stream_ptr()), "pio_gemm_f32")
    return c, rs


@pytest.mark.parametrize("m,n,k", [(4096, 256, 256), (100, 10, 131), (1, 7, 5), (300, 768, 256), (129, 130, 17),
                                   (2048, 10003, 64)])
def test_gemm_forward_layout_and_epilogues(N, m, n, k):
    torch.manual_seed(3)
    x = torch.randn(m, k)
    w = torch.randn(n, k) / math.sqrt(k)
    bias = torch.randn(n)
    res = torch.randn(m, n)
    xd, wd, bd, rd = (t.to(DEV) for t in (x, w, bias, res))
    ref = x.double() @ w.double().t()
    c, _ = _gemm(N, xd, wd, m, n, k, True, True)
    assert_rel(c, ref, 1e-5, "plain")
    c, _ = _gemm(N, xd, wd, m, n, k, True, True, bias=bd, residual=rd)
    assert_rel(c, ref + bias.double() + res.double(), 1e-5, "bias+res")
    aux = torch.empty(m, n, device=DEV)
    c, _ = _gemm(N, xd, wd, m, n, k, True, True, bias=bd, aux=aux, act=1)
    pre = ref + bias.double()
    assert_rel(aux, pre, 1e-5, "preact")
    assert_rel(c, O.gelu_erf(pre), 1e-5, "gelu")
    # act=2: multiply by GELU'(aux)
    pre_t = pre.clone().requires_grad_(True)
    O.gelu_erf(pre_t).sum().backward()
    c, _ = _gemm(N, xd, wd, m, n, k, True, True, aux=aux, act=2)
    assert_rel(c, ref * pre_t.grad, 2e-5, "dgelu")
    # accumulate
    c0 = torch.ones(m, n, device=DEV)
    c, _ = _gemm(N, xd, wd, m, n, k, True, True, accumulate=1, c=c0)
    assert_rel(c, ref + 1, 1e-5, "accumulate")


@pytest.mark.parametrize("rows,nout,kin", [(4096, 256, 256), (32768, 512, 256), (1000, 10, 131), (77, 5, 3),
                                           (8192, 768, 256)])
def test_gemm_backward_layouts_and_splitk(N, rows, nout, kin):
    torch.manual_seed(4)
    dy = torch.randn(rows, nout)
    x = torch.randn(rows, kin)
    w = torch.randn(nout, kin)
    dyd, xd, wd = dy.to(DEV), x.to(DEV), w.to(DEV)
    # dx = dy W   (A k-major, B not)
    dx, _ = _gemm(N, dyd, wd, rows, kin, nout, True, False)
    assert_rel(dx, dy.double() @ w.double(), 1e-5, "dx")
    # dW = dy^T x with the bias gradient riding along (both operands m/n-major, split-K)
    dw, db = _gemm(N, dyd, xd, nout, kin, rows, False, False, rowsum=True)
    assert_rel(dw, dy.double().t() @ x.double(), 2e-5, "dW")
    assert_rel(db, dy.double().sum(0), 2e-5, "db")
    # deterministic: two runs bit-identical
    dw2, db2 = _gemm(N, dyd, xd, nout, kin, rows, False, False, rowsum=True)
    assert torch.equal(dw, dw2) and torch.equal(db, db2)


def _attn_ref(q, k, v, pad, heads):
    b, n, e = q.shape
    m = k.shape[1]
    dh = e // heads
    qh = q.view(b, n, heads, dh).transpose(1, 2)
    kh = k.view(b, m, heads, dh).transpose(1, 2)
    vh = v.view(b, m, heads, dh).transpose(1, 2)
    s = (qh * dh ** -0.5) @ kh.transpose(-1, -2)
    if pad is not None:
        s = s.masked_fill(pad[:, None, None, :], float("-inf"))
    p = torch.softmax(s, -1)
    return (p @ vh).transpose(1, 2).reshape(b, n, e), torch.logsumexp(s, -1)


@pytest.mark.parametrize("b,n,m,heads,dh,padded", [
    (2, 8, 12, 2, 8, True), (3, 64, 512, 4, 64, True), (2, 32, 784, 4, 32, False), (2, 33, 65, 1, 128, True),
    (2, 1, 64, 1, 256, False), (1, 5, 40, 1, 512, True), (2, 70, 70, 4, 16, False), (1, 16, 2048, 8, 64, True)])
def test_attention_core_fwd_bwd(N, b, n, m, heads, dh, padded):
    torch.manual_seed(5)
    e = heads * dh
    q, k, v = torch.randn(b, n, e), torch.randn(b, m, e), torch.randn(b, m, e)
    do = torch.randn(b, n, e)
    pad = None
    if padded:
        lens = torch.randint(1, m + 1, (b,))
        lens[0] = m
        pad = torch.arange(m)[None, :] >= lens[:, None]
    qr, kr, vr = (t.double().requires_grad_(True) for t in (q, k, v))
    o_ref, lse_ref = _attn_ref(qr, kr, vr, pad, heads)
    (o_ref * do.double()).sum().backward()
    L = N.lib()
    qd, kd, vd, dod = (t.to(DEV) for t in (q, k, v, do))
    padd = pad.to(DEV) if pad is not None else None
    o = torch.empty(b, n, e, device=DEV)
    lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = qd.data_ptr(), kd.data_ptr(), vd.data_ptr(), N.ptr(padd), o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * e, e, m * e, e, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    N.check(L.pio_attn_fwd(ctypes.byref(d), 0, N.stream_ptr()), "pio_attn_fwd")
    assert_rel(o, o_ref, 1e-5, "attn o")
    assert_rel(lse, lse_ref, 1e-5, "attn lse")
    bd = N.AttnBwdDesc()
    bd.fwd = d
    dq, dk, dv = torch.empty_like(qd), torch.empty_like(kd), torch.empty_like(vd)
    delta = torch.empty(b, heads, n, device=DEV)
    bd.d_o, bd.dq, bd.dk, bd.dv, bd.delta = dod.data_ptr(), dq.data_ptr(), dk.data_ptr(), dv.data_ptr(), delta.data_ptr()
    N.check(L.pio_attn_bwd(ctypes.byref(bd), 0, N.stream_ptr()), "pio_attn_bwd")
    assert_rel(dq, qr.grad, 2e-5, "dq")
    assert_rel(dk, kr.grad, 2e-5, "dk")
    assert_rel(dv, vr.grad, 2e-5, "dv")


def test_attention_all_keys_padded_gives_nan_like_reference(N):
    b, n, m, heads, dh = 2, 4, 6, 1, 8
    e = heads * dh
    q, k, v = (torch.randn(b, x, e, device=DEV) for x in (n, m, m))
    pad = torch.zeros(b, m, dtype=torch.bool, device=DEV)
    pad[1] = True
    o = torch.empty(b, n, e, device=DEV)
    lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = q.data_ptr(), k.data_ptr(), v.data_ptr(), pad.data_ptr(), o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * e, e, m * e, e, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    N.check(N.lib().pio_attn_fwd(ctypes.byref(d), 0, N.stream_ptr()))
    assert torch.isfinite(o[0]).all() and torch.isnan(o[1]).all()


@pytest.mark.parametrize("rows,classes", [(48, 50), (4096, 10003), (7, 2), (1, 1)])
def test_cross_entropy_fwd_bwd(N, rows, classes):
    from perceiver import cross_entropy
    torch.manual_seed(6)
    z = torch.randn(rows, classes) * 3
    y = torch.randint(0, classes, (rows,))
    y[torch.rand(rows) < 0.5] = -100
    y[0] = 0
    zr = z.double().requires_grad_(True)
    ref = O.cross_entropy(zr, y)
    (ref * 1.7).backward()
    zd = z.to(DEV).requires_grad_(True)
    got = cross_entropy(zd, y.to(DEV))
    (got * 1.7).backward()
    assert_rel(got, ref, 1e-5, "ce")
    assert_rel(zd.grad, zr.grad, 1e-5, "dlogits")


def test_bad_arguments_return_errors_not_crashes(N):
    L = N.lib()
    assert L.pio_image_encode(None, None, None, 1, 1, 1, 1, None) == 1
    assert L.pio_image_encode(None, None, None, 0, 1, 1, 1, None) == 0  # empty batch is a no-op
    assert b"null" in L.pio_last_error()
    d = N.AttnDesc()
    assert L.pio_attn_fwd(ctypes.byref(d), 0, None) == 1
    x = torch.zeros(4, 8, device=DEV)
    w = N.mlp_struct(ln_w=x, ln_b=x, w1=x, b1=x, w2=x, b2=x)
    # workspace too small must be reported, not overrun
    saved = torch.empty(L.pio_mlp_saved_bytes(4, 8, 0), dtype=torch.uint8, device=DEV)
    rc = L.pio_mlp_bwd(x.data_ptr(), x.data_ptr(), x.data_ptr(), ctypes.byref(w), saved.data_ptr(), ctypes.byref(w), 0, 4, 8, 1,
                       None, 0, None, 0, N.stream_ptr())
    assert rc == 3 and b"workspace" in L.pio_last_error()
    with pytest.raises(RuntimeError, match="pio_b200 error"):
        N.check(rc, "pio_mlp_bwd")
