"""PIO_MODE_BF16 (tcgen05 GEMMs, bf16 operands, fp32 accumulate / residual stream) against the
fp32 references: north_star tolerance rel 1e-2, measured as max-abs error over the tensor's
largest magnitude."""
import pytest
import torch

from conftest import load_golden
from gpu_util import BF16_TOL, assert_rel, bf16_grad_bound, rel_err
from oracle import perceiver_oracle as O
from test_dropin_api import build_from_meta

pytestmark = pytest.mark.gpu
DEV = "cuda"
# Logits, loss and the gradient as a whole (all parameters concatenated, relative L2) keep the north-star bf16
# tolerance of 1e-2.  Per tensor the bound comes from gpu_util.BF16_GRAD_BOUNDS (2e-2 for matrices, 3e-2 for
# LayerNorm / bias vectors, 5e-2 for the decoder-query LayerNorm terms — each measured and justified there).  Only the
# 8..16-channel golden fixtures keep a blanket 1e-1: with so few numbers per tensor a single bf16 rounding is a visible
# fraction of the largest entry (vectors of fewer than 32 numbers are held to the whole-gradient criterion only).
GRAD_TOL_TINY = 1e-1
GRAD_TOL = None  # = the per-tensor table
GLOBAL_GRAD_TOL = BF16_TOL


@pytest.fixture(autouse=True)
def bf16_mode():
    import perceiver
    perceiver.set_compute_mode("bf16")
    yield
    perceiver.set_compute_mode("fp32")


def _check_grads(model, ref_grads, tol, global_tol=None, min_numel=0):
    worst, worst_l2 = ("", 0.0), ("", 0.0)
    num = den = 0.0
    for k, p in model.named_parameters():
        g = ref_grads.get(k)
        if g is None or g.abs().max() < 1e-10:
            continue
        assert p.grad is not None, k
        assert torch.isfinite(p.grad).all(), k
        e = rel_err(p.grad, g)
        diff = (p.grad.detach().cpu().double() - g.double()).norm().item()
        l2 = diff / g.double().norm().item()
        num += diff ** 2
        den += g.double().norm().item() ** 2
        if tol is None:
            assert e <= bf16_grad_bound(k), f"grad {k}: rel err {e:.3e} > {bf16_grad_bound(k):.1e}"
        if e > worst[1] and g.numel() >= min_numel:
            worst = (k, e)
        if l2 > worst_l2[1]:
            worst_l2 = (k, l2)
    glob = (num / max(den, 1e-300)) ** 0.5
    print(f"gradient: global L2-rel {glob:.3e}; worst tensor max-rel {worst[1]:.3e} ({worst[0]}), "
          f"L2-rel {worst_l2[1]:.3e} ({worst_l2[0]})")
    assert tol is None or worst[1] <= tol, f"worst grad {worst}"
    if global_tol is not None:
        assert glob <= global_tol, f"global gradient L2-rel {glob}"


@pytest.mark.parametrize("name", ["img_tiny", "img_rgb_1layer", "mlm_tiny", "mlm_short_l", "seqclf_tiny"])
def test_golden_bf16(name):
    import perceiver
    meta, a, sd, grads = load_golden(name)
    model = build_from_meta(meta)
    model.load_state_dict(sd)
    model = model.to(DEV)
    if meta["kind"] == "image":
        logits = model(a["x"].to(DEV))
        loss = perceiver.cross_entropy(logits, a["y"].to(DEV))
    elif meta.get("clf_classes") is None:
        l = a["ids_masked"].shape[1]
        lat, pm = model[0](a["ids_masked"].to(DEV), a["pad"].to(DEV))
        logits = model[1](lat, pm, num_queries=l)
        loss = perceiver.cross_entropy(logits, a["labels"].to(DEV))
    else:
        logits = model(a["ids"].to(DEV), a["pad"].to(DEV))
        loss = perceiver.cross_entropy(logits, a["y"].to(DEV))
    loss.backward()
    assert_rel(logits, a["logits"], BF16_TOL, "logits")
    assert_rel(loss, a["loss"], BF16_TOL, "loss")
    _check_grads(model, grads, GRAD_TOL_TINY, global_tol=2 * GLOBAL_GRAD_TOL, min_numel=32)


def _vs_oracle(kind, model, x, pad, target, kw):
    import perceiver
    sd_cpu = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}
    ref_logits = O.perceiver_io(sd_cpu, x, pad, kind=kind, **kw)
    ref_loss = O.cross_entropy(ref_logits, target)
    ref_loss.backward()
    model = model.to(DEV)
    got = model(x.to(DEV), pad.to(DEV)) if pad is not None else model(x.to(DEV))
    loss = perceiver.cross_entropy(got, target.to(DEV))
    loss.backward()
    assert_rel(got, ref_logits, BF16_TOL, "logits")
    assert_rel(loss, ref_loss, BF16_TOL, "loss")
    _check_grads(model, {k: v.grad for k, v in sd_cpu.items() if v.grad is not None}, GRAD_TOL, global_tol=GLOBAL_GRAD_TOL)


def test_config1_mnist_shape_bf16():
    from perceiver import ClassificationOutputAdapter, ImageInputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO
    torch.manual_seed(0)
    enc = PerceiverEncoder(ImageInputAdapter((28, 28, 1), 32), (32, 128), 3, 4, 4, 3)
    dec = PerceiverDecoder(ClassificationOutputAdapter(10, num_output_channels=128), (32, 128), 1)
    x = torch.rand(16, 28, 28, 1) * 2 - 1
    y = torch.randint(0, 10, (16,))
    _vs_oracle("image", PerceiverIO(enc, dec), x, None, y,
               dict(num_layers=3, cross_heads=4, self_heads=4, sa_layers=3, dec_heads=1))


def test_config2_mlm_shape_bf16():
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter, TextOutputAdapter
    torch.manual_seed(0)
    enc = PerceiverEncoder(TextInputAdapter(10003, 512, 256), (64, 256), 3, 4, 4, 6)
    dec = PerceiverDecoder(TextOutputAdapter(10003, 512, 256), (64, 256), 4)
    b = 4
    ids = torch.randint(3, 10003, (b, 512))
    lens = torch.randint(128, 513, (b,))
    lens[0] = 512
    pad = torch.arange(512)[None, :] >= lens[:, None]
    ids[pad] = 0
    labels = torch.full((b, 512), -100, dtype=torch.long)
    sel = (torch.rand(b, 512) < 0.15) & ~pad
    labels[sel] = ids[sel]
    _vs_oracle("text", PerceiverIO(enc, dec), ids, pad, labels,
               dict(num_layers=3, cross_heads=4, self_heads=4, sa_layers=6, dec_heads=4))


def test_weight_shadow_follows_parameter_updates():
    """The bf16 shadows are keyed on the parameter version: an optimizer step must be visible."""
    from perceiver import mlp
    torch.manual_seed(1)
    ff = mlp(32).to(DEV)
    x = torch.randn(4, 8, 32, device=DEV)
    y0 = ff(x)
    with torch.no_grad():
        ff[1].weight.mul_(2.0)
    y1 = ff(x)
    assert rel_err(y1, y0) > 1e-2
    sd = {"module." + k: v.cpu() for k, v in ff.state_dict().items()}
    assert_rel(y1, O.mlp_residual(sd, "", x.cpu()) - x.cpu(), BF16_TOL, "after update")


@pytest.mark.parametrize("rows,c,classes,p_ignore", [(300, 64, 517, 0.8), (4096, 256, 10003, 0.8), (9000, 64, 517, 0.0),
                                                     (9001, 72, 1001, 0.05), (257, 131, 40, 0.5)])
def test_fused_linear_cross_entropy_matches_oracle(rows, c, classes, p_ignore):
    """pio_linear_ce_fwd_bwd (valid rows compacted on the device, chunked, logits never stored) against the fp64
    oracle of linear + CE; rows past one chunk (8192), counts off the 64/128-row tile grid, no ignored rows."""
    import perceiver
    torch.manual_seed(7)
    x = torch.randn(rows, c)
    w = torch.randn(classes, c) * 0.1
    b = torch.randn(classes) * 0.1
    y = torch.randint(0, classes, (rows,))
    y[torch.rand(rows) < p_ignore] = -100
    y[0] = 3
    xr, wr, br = (t.double().requires_grad_(True) for t in (x, w, b))
    ref = O.cross_entropy(O.linear(xr, wr, br), y)
    (ref * 1.5).backward()
    xd, wd, bd = (t.to(DEV).requires_grad_(True) for t in (x, w, b))
    loss = perceiver.linear_cross_entropy(xd, wd, bd, y.to(DEV))
    (loss * 1.5).backward()
    assert_rel(loss, ref, BF16_TOL, "loss")
    assert_rel(xd.grad, xr.grad, 2e-2, "dx")
    assert_rel(wd.grad, wr.grad, 2e-2, "dw")
    assert_rel(bd.grad, br.grad, 2e-2, "db")
    assert not xd.grad[(y == -100).to(DEV)].any()  # ignored rows: exactly zero


def test_fused_linear_cross_entropy_graph_replay_with_other_label_counts():
    """The number of labelled rows never reaches the host: a CUDA graph captured on one batch gives the eager
    result on batches with fewer, more, or no ignored labels."""
    from perceiver import _native as N
    L = N.lib()
    rows, c, classes = 3000, 64, 517
    torch.manual_seed(11)
    x = torch.randn(rows, c, device=DEV)
    w = (torch.randn(classes, c, device=DEV) * 0.1)
    w16 = w.to(torch.bfloat16).contiguous()
    b = torch.randn(classes, device=DEV) * 0.1
    labels = torch.empty(rows, dtype=torch.int64, device=DEV)
    ws = torch.empty(L.pio_linear_ce_ws_bytes(rows, c, classes, 1), dtype=torch.uint8, device=DEV)

    def make(p_ignore, seed):
        g = torch.Generator().manual_seed(seed)
        y = torch.randint(0, classes, (rows,), generator=g)
        y[torch.rand(rows, generator=g) < p_ignore] = -100
        y[5] = 1
        return y.to(DEV)

    def run(out3, dx, dw, db):
        N.check(L.pio_linear_ce_fwd_bwd(x.data_ptr(), w16.data_ptr(), b.data_ptr(), labels.data_ptr(), out3.data_ptr(),
                                        dx.data_ptr(), dw.data_ptr(), db.data_ptr(), rows, c, classes, -100, 1, ws.data_ptr(),
                                        ws.numel(), N.stream_ptr(x.device)), "pio_linear_ce_fwd_bwd")

    outs = [torch.empty(3, device=DEV), torch.empty(rows, c, device=DEV), torch.empty(classes, c, device=DEV),
            torch.empty(classes, device=DEV)]
    labels.copy_(make(0.85, 1))
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        run(*outs)
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    with torch.cuda.graph(graph):
        run(*outs)
    for p_ignore, seed in [(0.85, 1), (0.3, 2), (0.0, 3), (0.97, 4), (0.5, 5)]:
        labels.copy_(make(p_ignore, seed))
        graph.replay()
        torch.cuda.synchronize()
        got = [t.clone() for t in outs]
        eager = [torch.empty_like(t) for t in outs]
        run(*eager)
        torch.cuda.synchronize()
        assert got[0][1].item() == (labels != -100).sum().item()
        assert_rel(got[0][0], eager[0][0], 1e-5, f"loss p={p_ignore}")
        assert torch.equal(got[1], eager[1]), f"dx p={p_ignore}"
        assert_rel(got[2], eager[2], 1e-4, f"dw p={p_ignore}")  # atomics: summation order differs
        assert_rel(got[3], eager[3], 1e-4, f"db p={p_ignore}")


def test_mlm_return_loss_equals_logits_then_cross_entropy():
    """PerceiverMLM(..., return_loss=True) == cross_entropy(PerceiverMLM(...)[0], labels), values and grads."""
    import perceiver
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    torch.manual_seed(8)
    enc = PerceiverEncoder(TextInputAdapter(1003, 128, 64), (32, 64), 2, 1, 1, 2)
    dec = PerceiverDecoder(TextOutputAdapter(1003, 128, 64), (32, 64), 1)
    model = PerceiverMLM(enc, dec, TextMasking(1003, 1, 2, 3)).to(DEV)
    ids = torch.randint(3, 1003, (6, 100), device=DEV)
    pad = torch.arange(100, device=DEV)[None, :] >= torch.tensor([100, 90, 80, 70, 60, 50], device=DEV)[:, None]
    g1 = torch.Generator(device=DEV).manual_seed(1)
    torch.manual_seed(5)
    logits, labels = model(ids.clone(), pad)
    loss_a = perceiver.cross_entropy(logits, labels)
    loss_a.backward()
    ga = {k: p.grad.clone() for k, p in model.named_parameters()}
    model.zero_grad(set_to_none=True)
    torch.manual_seed(5)
    loss_b, labels_b = model(ids.clone(), pad, return_loss=True)
    loss_b.backward()
    assert torch.equal(labels, labels_b)
    assert_rel(loss_b, loss_a, 2e-3, "loss")
    num = sum(((p.grad - ga[k]).double().norm() ** 2).item() for k, p in model.named_parameters())
    den = sum((ga[k].double().norm() ** 2).item() for k in ga)
    assert (num / den) ** 0.5 < 1e-2


def test_config4_long_input_shape_bf16():
    """BASELINE configs[3] scaled down (M=2048, latent (256,512), 8 cross heads of 64 -> tcgen05 flash
    kernels; 4 self heads of 128 and a 1-head decoder of 512 -> fp32-arithmetic kernels on bf16 operands)."""
    from perceiver import ClassificationOutputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter
    torch.manual_seed(2)
    enc = PerceiverEncoder(TextInputAdapter(10003, 2048, 512), (256, 512), 2, 8, 4, 2)
    dec = PerceiverDecoder(ClassificationOutputAdapter(2, num_output_channels=512), (256, 512), 1)
    b = 2
    ids = torch.randint(3, 10003, (b, 2048))
    pad = torch.arange(2048)[None, :] >= torch.tensor([2048, 1500])[:, None]
    ids[pad] = 0
    y = torch.randint(0, 2, (b,))
    _vs_oracle("text", PerceiverIO(enc, dec), ids, pad, y,
               dict(num_layers=2, cross_heads=8, self_heads=4, sa_layers=2, dec_heads=1))


def test_bf16_shadows_follow_fused_optimizer_steps():
    """torch.optim.AdamW(fused=True) updates parameters WITHOUT moving their version counters, so the bf16
    weight shadows must be re-cast on every training forward (one pio_cast_bf16_multi launch from the second
    forward on), in place, and must equal the current fp32 weights rounded to bf16."""
    import perceiver
    from perceiver import functional as F
    from perceiver import _native as N
    perceiver.set_compute_mode("bf16")
    try:
        torch.manual_seed(0)
        meta = dict(vocab=503, seq=64, latent=(16, 64), layers=2, sa=2)
        from perceiver import (PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking,
                               TextOutputAdapter)
        enc = PerceiverEncoder(TextInputAdapter(meta["vocab"], meta["seq"], 64), meta["latent"], meta["layers"], 4, 4, meta["sa"])
        dec = PerceiverDecoder(TextOutputAdapter(meta["vocab"], meta["seq"], num_output_channels=64), meta["latent"], 4)
        model = PerceiverMLM(enc, dec, TextMasking(meta["vocab"], 1, 2, 3)).cuda()
        opt = torch.optim.AdamW(model.parameters(), lr=1e-2, fused=True)
        ids = torch.randint(3, meta["vocab"], (8, meta["seq"]), device="cuda")
        pad = torch.zeros(8, meta["seq"], dtype=torch.bool, device="cuda")
        ptrs = None
        for it in range(4):
            before = N.lib().pio_launch_count()
            loss, _ = model(ids.clone(), pad, return_loss=True)
            shadowed = [p for p in model.parameters() if getattr(p, "_pio_w16", None) is not None]
            assert len(shadowed) >= 10
            for p in shadowed:
                cols = p.shape[1]
                assert torch.equal(p._pio_w16[2][:, :cols], p.detach().to(torch.bfloat16)), f"stale shadow at step {it}"
                assert not p._pio_w16[2][:, cols:].any()
            now = [p._pio_w16[2].data_ptr() for p in shadowed]
            if ptrs is not None:
                assert now == ptrs  # refreshed in place
            ptrs = now
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
        assert model in F._TABLES and len(F._TABLES[model][2]) == len(shadowed)
        # eval after the last optimizer step must not reuse the last training forward's shadows
        model.eval()
        with torch.no_grad():
            model(ids.clone(), pad)
            for p in shadowed:
                assert torch.equal(p._pio_w16[2][:, :p.shape[1]], p.detach().to(torch.bfloat16))
    finally:
        perceiver.set_compute_mode("fp32")
