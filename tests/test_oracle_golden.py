"""Pin oracle/perceiver_oracle.py to the fixtures produced by the unmodified reference
(oracle/make_golden.py).  CPU only."""
import math

import pytest
import torch

from conftest import load_golden
from oracle import perceiver_oracle as O
from oracle import reference_loader

TOL = dict(rtol=1e-5, atol=1e-6)


def _oracle_loss(meta, arrays, sd):
    sd = {k: v.clone().requires_grad_(v.is_floating_point()) for k, v in sd.items()}
    h = meta["heads"]
    kw = dict(num_layers=meta["layers"], cross_heads=h[0], self_heads=h[1], sa_layers=meta["sa"], dec_heads=h[2])
    if meta["kind"] == "image":
        logits = O.perceiver_io(sd, arrays["x"], None, kind="image", **kw)
        loss = O.cross_entropy(logits, arrays["y"])
    elif meta.get("clf_classes") is None:
        l = arrays["ids_masked"].shape[1]
        logits = O.perceiver_io(sd, arrays["ids_masked"], arrays["pad"], kind="text", **kw)[:, :l]
        loss = O.cross_entropy(logits, arrays["labels"])
    else:
        logits = O.perceiver_io(sd, arrays["ids"], arrays["pad"], kind="text", **kw)
        loss = O.cross_entropy(logits, arrays["y"])
    loss.backward()
    return logits, loss, sd


@pytest.mark.parametrize("name", ["img_tiny", "img_rgb_1layer", "mlm_tiny", "mlm_short_l", "seqclf_tiny"])
def test_oracle_matches_reference_forward_and_grads(name):
    meta, arrays, sd, grads = load_golden(name)
    logits, loss, sd_g = _oracle_loss(meta, arrays, sd)
    torch.testing.assert_close(logits.detach(), arrays["logits"], **TOL)
    torch.testing.assert_close(loss.detach(), arrays["loss"], **TOL)
    assert set(grads) <= set(sd_g)
    for k, g in grads.items():
        got = sd_g[k].grad
        got = torch.zeros_like(g) if got is None else got
        torch.testing.assert_close(got, g, rtol=1e-4, atol=1e-6, msg=lambda m, k=k: f"{k}: {m}")


def test_oracle_masking_bit_exact():
    meta, a, _, _ = load_golden("masking_10003")
    g = torch.Generator()
    g.set_state(a["rng_state"])
    xm, labels = O.text_masking(a["ids"].clone(), a["pad"], vocab_size=meta["vocab"], unk_token_id=1,
                                mask_token_id=2, num_special_tokens=3, generator=g)
    assert torch.equal(xm, a["ids_masked"])
    assert torch.equal(labels, a["labels"])
    sel = labels != -100
    assert 0.10 < sel.float().mean() / (~a["pad"] & (a["ids"] != 1)).float().mean() < 0.20
    assert not (sel & (a["pad"] | (a["ids"] == 1))).any()


@pytest.mark.parametrize("name", ["fourier_6x5_f3", "fourier_3d"])
def test_oracle_fourier_table(name):
    meta, a, _, _ = load_golden(name)
    shape = meta["image_shape"]
    enc = O.fourier_position_encoding(shape[:-1], meta["bands"])
    torch.testing.assert_close(enc, a["position_encoding"], rtol=0, atol=1e-6)
    d = len(shape) - 1
    assert torch.equal(enc[:, :d], a["position_encoding"][:, :d])
    out = O.image_input_adapter({"position_encoding": a["position_encoding"]}, "", a["x"])
    assert torch.equal(out, a["out"])


def test_oracle_fourier_mnist_rows():
    meta, a, _, _ = load_golden("fourier_28x28_f32_rows")
    enc = O.fourier_position_encoding((28, 28), 32)
    assert enc.shape == (784, 130)
    torch.testing.assert_close(enc[::meta["stride"]], a["rows"], rtol=0, atol=2e-6)
    assert math.isclose(float(enc.double().sum()), float(a["sum"]), abs_tol=1e-2)


@pytest.mark.skipif(not reference_loader.available(), reason="reference tree only exists in the build container")
def test_oracle_against_live_reference_cfg1_shape():
    """MNIST-shaped config (BASELINE configs[0]) at reduced batch, against the live reference."""
    import subprocess, sys, os
    code = r'''
import sys, torch
sys.path.insert(0, %r)
from oracle import reference_loader, perceiver_oracle as O
M, A = reference_loader.load()
torch.manual_seed(0)
enc = M.PerceiverEncoder(A.ImageInputAdapter((28, 28, 1), 32), (32, 128), 3, 4, 4, 3)
dec = M.PerceiverDecoder(A.ClassificationOutputAdapter(10, num_output_channels=128), (32, 128), 1)
model = M.PerceiverIO(enc, dec)
x = torch.rand(4, 28, 28, 1) * 2 - 1
ref = model(x)
got = O.perceiver_io(dict(model.state_dict()), x, None, kind="image", num_layers=3, cross_heads=4,
                     self_heads=4, sa_layers=3, dec_heads=1)
assert (ref - got).abs().max().item() < 1e-5, (ref - got).abs().max().item()
print("OK")
''' % os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    out = subprocess.run([sys.executable, "-c", code], capture_output=True, text=True, timeout=300)
    assert out.returncode == 0 and "OK" in out.stdout, out.stderr[-2000:]
