"""End-to-end parity of the drop-in modules (CUDA path through the C ABI) against
 (a) the committed golden fixtures produced by the unmodified reference, and
 (b) the CPU oracle at BASELINE.json-shaped configs, plus size-independent properties at full size."""
import pytest
import torch

from conftest import load_golden
from gpu_util import FP32_TOL, assert_rel, rel_err
from oracle import perceiver_oracle as O
from test_dropin_api import build_from_meta

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _run_golden(name):
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
        # decoding every query then slicing (the reference's order) gives the same logits
        full = model[1](lat.detach(), pm)[:, :l]
        assert_rel(full, a["logits"], FP32_TOL, "full-K decode")
    else:
        logits = model(a["ids"].to(DEV), a["pad"].to(DEV))
        loss = perceiver.cross_entropy(logits, a["y"].to(DEV))
    loss.backward()
    return model, logits, loss, a, grads


@pytest.mark.parametrize("name", ["img_tiny", "img_rgb_1layer", "mlm_tiny", "mlm_short_l", "seqclf_tiny"])
def test_golden_forward_and_all_gradients(name):
    model, logits, loss, a, grads = _run_golden(name)
    assert logits.shape == a["logits"].shape
    assert_rel(logits, a["logits"], FP32_TOL, "logits")
    assert_rel(loss, a["loss"], FP32_TOL, "loss")
    got = dict(model.named_parameters())
    assert set(got) == set(grads)
    for k, g in grads.items():
        gg = got[k].grad
        if gg is None:
            assert g.abs().max() == 0, f"{k}: missing grad"
            continue
        if g.abs().max() < 1e-10:
            assert gg.abs().max().item() < 1e-7, k
            continue
        assert_rel(gg, g, FP32_TOL, k)


def _oracle_vs_cuda(kind, model, x, pad, target, kw, is_mlm=False):
    import perceiver
    sd_cpu = {k: v.detach().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}
    ref_logits = O.perceiver_io(sd_cpu, x, pad, kind=kind, **kw)
    ref_loss = O.cross_entropy(ref_logits, target)
    ref_loss.backward()
    model = model.to(DEV)
    got_logits = model(x.to(DEV), pad.to(DEV)) if pad is not None else model(x.to(DEV))
    got_loss = perceiver.cross_entropy(got_logits, target.to(DEV))
    got_loss.backward()
    assert_rel(got_logits, ref_logits, FP32_TOL, "logits")
    assert_rel(got_loss, ref_loss, FP32_TOL, "loss")
    worst = ("", 0.0)
    for k, p in model.named_parameters():
        rg = sd_cpu[k].grad
        if rg is None or rg.abs().max() < 1e-12:
            continue
        e = rel_err(p.grad, rg)
        if e > worst[1]:
            worst = (k, e)
    assert worst[1] <= FP32_TOL, f"worst grad {worst}"


def test_config1_mnist_shape_vs_oracle():
    """BASELINE configs[0]: ImageInputAdapter 28x28x1, 32 bands, latent (32,128), 3 layers x 3 SA,
    heads 4/4/1 (reduced batch so the CPU oracle finishes in seconds)."""
    from perceiver import ClassificationOutputAdapter, ImageInputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO
    torch.manual_seed(0)
    enc = PerceiverEncoder(ImageInputAdapter((28, 28, 1), 32), (32, 128), 3, 4, 4, 3)
    dec = PerceiverDecoder(ClassificationOutputAdapter(10, num_output_channels=128), (32, 128), 1)
    model = PerceiverIO(enc, dec)
    x = torch.rand(16, 28, 28, 1) * 2 - 1
    y = torch.randint(0, 10, (16,))
    _oracle_vs_cuda("image", model, x, None, y, dict(num_layers=3, cross_heads=4, self_heads=4, sa_layers=3, dec_heads=1))


def test_config2_mlm_shape_vs_oracle():
    """BASELINE configs[1]: vocab 10003, seq 512, latent (64,256), 3 x 6 SA, heads 4/4/4, IMDB-shaped
    padding + UNK tokens (reduced batch)."""
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter, TextOutputAdapter
    torch.manual_seed(0)
    enc = PerceiverEncoder(TextInputAdapter(10003, 512, 256), (64, 256), 3, 4, 4, 6)
    dec = PerceiverDecoder(TextOutputAdapter(10003, 512, 256), (64, 256), 4)
    model = PerceiverIO(enc, dec)
    b = 4
    ids = torch.randint(3, 10003, (b, 512))
    lens = torch.randint(128, 513, (b,))
    lens[0] = 512
    pad = torch.arange(512)[None, :] >= lens[:, None]
    ids[torch.rand(b, 512) < 0.02] = 1
    ids[pad] = 0
    labels = torch.full((b, 512), -100, dtype=torch.long)
    sel = (torch.rand(b, 512) < 0.15) & ~pad
    labels[sel] = ids[sel]
    _oracle_vs_cuda("text", model, ids, pad, labels,
                    dict(num_layers=3, cross_heads=4, self_heads=4, sa_layers=6, dec_heads=4))


def test_config3_seqclf_shape_vs_oracle():
    """BASELINE configs[2]: MLM encoder + ClassificationOutputAdapter, decoder 1 head (dh = 256)."""
    from perceiver import ClassificationOutputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter
    torch.manual_seed(1)
    enc = PerceiverEncoder(TextInputAdapter(10003, 512, 256), (64, 256), 3, 4, 4, 6)
    dec = PerceiverDecoder(ClassificationOutputAdapter(2, num_output_channels=256), (64, 256), 1)
    model = PerceiverIO(enc, dec)
    b = 6
    ids = torch.randint(3, 10003, (b, 512))
    lens = torch.randint(128, 513, (b,))
    pad = torch.arange(512)[None, :] >= lens[:, None]
    ids[pad] = 0
    y = torch.randint(0, 2, (b,))
    _oracle_vs_cuda("text", model, ids, pad, y, dict(num_layers=3, cross_heads=4, self_heads=4, sa_layers=6, dec_heads=1))


def test_config4_long_input_shape_vs_oracle():
    """BASELINE configs[3] scaled down in batch and M (M=2048, latent (256,512), 8 cross heads, 1 layer
    of 2 SA so the CPU oracle stays in seconds): exercises dh=64 cross / dh=128 self / dh=512 decoder."""
    from perceiver import ClassificationOutputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO, TextInputAdapter
    torch.manual_seed(2)
    enc = PerceiverEncoder(TextInputAdapter(10003, 2048, 512), (256, 512), 2, 8, 4, 2)
    dec = PerceiverDecoder(ClassificationOutputAdapter(2, num_output_channels=512), (256, 512), 1)
    model = PerceiverIO(enc, dec)
    b = 2
    ids = torch.randint(3, 10003, (b, 2048))
    pad = torch.arange(2048)[None, :] >= torch.tensor([2048, 1500])[:, None]
    ids[pad] = 0
    y = torch.randint(0, 2, (b,))
    _oracle_vs_cuda("text", model, ids, pad, y, dict(num_layers=2, cross_heads=8, self_heads=4, sa_layers=2, dec_heads=1))


def _mlm_model():
    from perceiver import (PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking,
                           TextOutputAdapter)
    enc = PerceiverEncoder(TextInputAdapter(10003, 512, 256), (64, 256), 3, 4, 4, 6)
    dec = PerceiverDecoder(TextOutputAdapter(10003, 512, 256), (64, 256), 4)
    return PerceiverMLM(enc, dec, TextMasking(10003, 1, 2, 3))


def test_full_size_mlm_properties():
    """Config 2 at its full size (batch 64, seq 512) — properties that need no oracle run:
    (1) samples are independent: a sub-batch alone gives the same logits as inside the batch;
    (2) token ids under the pad mask do not influence any unpadded output;
    (3) gradients are additive over the batch (sum-loss grads of two halves = grads of the whole)."""
    import perceiver
    torch.manual_seed(0)
    model = _mlm_model().to(DEV)
    b = 64
    ids = torch.randint(3, 10003, (b, 512), device=DEV)
    lens = torch.randint(128, 513, (b,), device=DEV)
    lens[0] = 512
    pad = torch.arange(512, device=DEV)[None, :] >= lens[:, None]
    ids[pad] = 0
    with torch.no_grad():
        full, _ = model(ids.clone(), pad, masking=False)
        assert full.shape == (b, 512, 10003) and torch.isfinite(full).all()
        part, _ = model(ids[:3].clone(), pad[:3], masking=False)
        assert_rel(part, full[:3], 1e-5, "batch independence")
        ids2 = ids.clone()
        ids2[pad] = torch.randint(3, 10003, (int(pad.sum()),), device=DEV)
        other, _ = model(ids2, pad, masking=False)
        assert_rel(other, full, 1e-6, "padded ids are ignored")

    def grads(sl):
        model.zero_grad(set_to_none=True)
        logits, labels = model(ids[sl].clone(), pad[sl], masking=False)
        tgt = torch.where(pad[sl], torch.full_like(ids[sl], -100), ids[sl])
        loss = perceiver.cross_entropy(logits, tgt)
        n_valid = (tgt != -100).sum()
        (loss * n_valid).backward()  # sum-loss so halves add up
        return {k: p.grad.clone() for k, p in model.named_parameters()}
    g_all = grads(slice(0, 8))
    g_a = grads(slice(0, 4))
    g_b = grads(slice(4, 8))
    for k in g_all:
        assert_rel(g_a[k] + g_b[k], g_all[k], 1e-4, f"additivity {k}")


def test_mlm_module_wiring_and_masking():
    """PerceiverMLM.forward: masks in place, returns (logits[:, :l], labels); shorter-than-max
    sequences only decode l queries."""
    torch.manual_seed(3)
    model = _mlm_model().to(DEV)
    ids = torch.randint(3, 10003, (4, 300), device=DEV)
    pad = torch.zeros(4, 300, dtype=torch.bool, device=DEV)
    orig = ids.clone()
    logits, labels = model(ids, pad)
    assert logits.shape == (4, 300, 10003) and labels.shape == (4, 300)
    changed = ids != orig
    assert changed.any() and ((labels != -100) | ~changed).all()
    logits2, none = model(orig.clone(), pad, masking=False)
    assert none is None and logits2.shape == (4, 300, 10003)


def test_layer_n_kv_reuse_matches_per_repeat_projection():
    """The shared layer_n projects K/V once; recomputing it per repeat (the reference's order of
    operations) must give the same latents and the same accumulated weight gradients."""
    from perceiver import PerceiverEncoder, TextInputAdapter
    torch.manual_seed(4)
    enc = PerceiverEncoder(TextInputAdapter(100, 32, 32), (8, 32), 4, 2, 2, 1).to(DEV)
    ids = torch.randint(3, 100, (3, 32), device=DEV)
    lat, _ = enc(ids)
    lat.square().sum().backward()
    g1 = {k: p.grad.clone() for k, p in enc.named_parameters()}
    enc.zero_grad(set_to_none=True)
    x = enc.input_adapter(ids)
    z = enc.latent.unsqueeze(0).expand(3, -1, -1).contiguous()
    z = enc._apply_layer(enc.layer_1, z, x, None, None)
    for _ in range(3):
        z = enc._apply_layer(enc.layer_n, z, x, None, None)  # kv=None -> projected again each time
    assert_rel(z, lat, 1e-6, "latents")
    z.square().sum().backward()
    for k, p in enc.named_parameters():
        assert_rel(p.grad, g1[k], 1e-5, k)


def test_standalone_modules_without_residual():
    from perceiver import CrossAttention, SelfAttention, mlp
    torch.manual_seed(5)
    x = torch.randn(2, 6, 16)
    kvx = torch.randn(2, 9, 12)
    ca, sa, ff = CrossAttention(16, 12, 2, 0.0), SelfAttention(16, 4, 0.0), mlp(16)
    sd = {"module." + k: v for k, v in ca.state_dict().items()}
    ref = O.cross_attention_residual(sd, "", x, kvx, 2, None) - x
    assert_rel(ca.to(DEV)(x.to(DEV), kvx.to(DEV)), ref, 1e-5, "CrossAttention")
    sd = {"module." + k: v for k, v in sa.state_dict().items()}
    assert_rel(sa.to(DEV)(x.to(DEV)), O.self_attention_residual(sd, "", x, 4) - x, 1e-5, "SelfAttention")
    sd = {"module." + k: v for k, v in ff.state_dict().items()}
    assert_rel(ff.to(DEV)(x.to(DEV)), O.mlp_residual(sd, "", x) - x, 1e-5, "mlp")


def test_attn_mask_is_rejected_not_ignored():
    from perceiver import PerceiverEncoder, TextInputAdapter
    enc = PerceiverEncoder(TextInputAdapter(50, 8, 16), (4, 16), 1, 2, 2, 1).to(DEV)
    ids = torch.randint(3, 50, (2, 8), device=DEV)
    with pytest.raises(NotImplementedError):
        enc(ids, None, torch.zeros(4, 8, device=DEV))
