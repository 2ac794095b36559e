"""Host-side logic of the drop-in classes: constructor signatures, state_dict keys/shapes equal to
the reference's (taken from the golden fixtures), checkpoint loading, init statistics, error
behaviour.  No GPU."""
import inspect

import pytest
import torch

from conftest import load_golden

import perceiver
from perceiver import (ClassificationOutputAdapter, ImageInputAdapter, PerceiverDecoder, PerceiverEncoder, PerceiverIO,
                       PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter)


def build_from_meta(meta):
    """Same construction order as oracle/make_golden.py so the RNG stream lines up."""
    h = meta["heads"]
    latent = tuple(meta["latent"])
    if meta["kind"] == "image":
        inp = ImageInputAdapter(tuple(meta["image_shape"]), meta["bands"])
    else:
        inp = TextInputAdapter(meta["vocab"], meta["seq"], meta["C"])
    enc = PerceiverEncoder(inp, latent, meta["layers"], num_cross_attention_heads=h[0], num_self_attention_heads=h[1],
                           num_self_attention_layers_per_block=meta["sa"])
    if meta["kind"] == "image":
        out = ClassificationOutputAdapter(meta["classes"], num_output_channels=latent[1])
    elif meta.get("clf_classes") is None:
        out = TextOutputAdapter(meta["vocab"], meta["seq"], num_output_channels=latent[1])
    else:
        out = ClassificationOutputAdapter(meta["clf_classes"], num_output_channels=latent[1])
    dec = PerceiverDecoder(out, latent, num_cross_attention_heads=h[2])
    return PerceiverIO(enc, dec)


@pytest.mark.parametrize("name", ["img_tiny", "img_rgb_1layer", "mlm_tiny", "seqclf_tiny"])
def test_state_dict_keys_and_shapes_match_reference(name):
    meta, _, sd, _ = load_golden(name)
    model = build_from_meta(meta)
    mine = model.state_dict()
    assert list(mine.keys()) == list(sd.keys()) or set(mine.keys()) == set(sd.keys())
    assert set(mine.keys()) == set(sd.keys())
    for k, v in sd.items():
        assert tuple(mine[k].shape) == tuple(v.shape), k
    model.load_state_dict(sd, strict=True)  # a reference checkpoint loads as is
    assert set(k for k, _ in model.named_parameters()) == set(k for k in sd if "position_encoding" not in k)


def test_fourier_buffer_equals_reference():
    _, a, _, _ = load_golden("fourier_6x5_f3")
    ad = ImageInputAdapter((6, 5, 1), 3)
    assert torch.equal(ad.position_encoding, a["position_encoding"])
    assert ad.num_input_channels == 1 + 2 * (2 * 3 + 1)
    _, a3, _, _ = load_golden("fourier_3d")
    assert torch.equal(ImageInputAdapter((3, 4, 2, 2), 2).position_encoding, a3["position_encoding"])


def test_constructor_signatures_match_reference():
    def names(f):
        return list(inspect.signature(f).parameters)
    assert names(PerceiverEncoder.__init__) == ["self", "input_adapter", "latent_shape", "num_layers",
                                                "num_cross_attention_heads", "num_self_attention_heads",
                                                "num_self_attention_layers_per_block", "dropout"]
    assert inspect.signature(PerceiverEncoder.__init__).parameters["num_self_attention_layers_per_block"].default == 2
    assert names(PerceiverDecoder.__init__) == ["self", "output_adapter", "latent_shape", "num_cross_attention_heads",
                                                "dropout"]
    assert names(TextMasking.__init__) == ["self", "vocab_size", "unk_token_id", "mask_token_id", "num_special_tokens",
                                           "mask_p"]
    assert names(ImageInputAdapter.__init__) == ["self", "image_shape", "num_frequency_bands"]
    assert names(TextInputAdapter.__init__) == ["self", "vocab_size", "max_seq_len", "num_input_channels"]
    assert names(ClassificationOutputAdapter.__init__) == ["self", "num_classes", "num_outputs", "num_output_channels"]
    assert names(TextOutputAdapter.__init__) == ["self", "vocab_size", "max_seq_len", "num_output_channels"]
    assert names(PerceiverMLM.__init__) == ["self", "encoder", "decoder", "masking"]
    assert names(PerceiverEncoder.forward)[:4] == ["self", "x", "pad_mask", "attn_mask"]


def test_init_follows_reference_distributions():
    torch.manual_seed(0)
    enc = PerceiverEncoder(TextInputAdapter(1000, 64, 32), (16, 32), 2)
    dec = PerceiverDecoder(TextOutputAdapter(1000, 64, 32), (16, 32))
    assert enc.latent.abs().max() <= 2.0 and 0.01 < enc.latent.std() < 0.03
    assert dec.output.shape == (64, 32) and 0.01 < dec.output.std() < 0.03
    emb = enc.input_adapter.text_embedding.weight
    assert emb.min() >= -0.1 and emb.max() <= 0.1
    assert enc.input_adapter.pos_encoding.abs().max() <= 0.5
    a = enc.layer_1[0][0].module.attention.attention
    assert torch.all(a.in_proj_bias == 0) and torch.all(a.out_proj.bias == 0)
    bound = (6.0 / (32 + 3 * 32)) ** 0.5  # xavier_uniform on (3E, E)
    assert a.in_proj_weight.abs().max() <= bound + 1e-6
    assert enc.output_shape if hasattr(enc, "output_shape") else True
    assert dec.output_adapter.output_shape == (64, 32)


def test_same_seed_gives_reference_identical_init():
    """Constructing with the seed make_golden.py used reproduces the reference's un-perturbed
    initial latent (the fixture's latent is untouched by _perturb)."""
    meta, _, sd, _ = load_golden("img_rgb_1layer")
    torch.manual_seed(1)
    model = build_from_meta(meta)
    assert torch.equal(model.state_dict()["0.latent"], sd["0.latent"])
    assert torch.equal(model.state_dict()["1.output"], sd["1.output"])
    k = "0.layer_1.1.0.1.module.1.weight"
    assert torch.equal(model.state_dict()[k], sd[k])


def test_errors_match_reference_behaviour():
    ad = ImageInputAdapter((6, 5, 1), 3)
    with pytest.raises(ValueError, match="different from required shape"):
        ad(torch.zeros(2, 5, 6, 1))
    dec = PerceiverDecoder(ClassificationOutputAdapter(3, num_output_channels=8), (4, 8), 1)
    with pytest.raises(ValueError, match="Latent shape"):
        dec(torch.zeros(2, 5, 8), None)
    with pytest.raises(AssertionError):
        PerceiverEncoder(TextInputAdapter(10, 4, 6), (4, 6), 1, num_cross_attention_heads=4)


def test_no_cpu_fallback():
    ad = ImageInputAdapter((6, 5, 1), 3)
    with pytest.raises(RuntimeError, match="CUDA"):
        ad(torch.zeros(2, 6, 5, 1))
    enc = PerceiverEncoder(TextInputAdapter(10, 4, 8), (4, 8), 1, 2, 2, 1)
    with pytest.raises(RuntimeError, match="CUDA"):
        enc(torch.zeros(2, 4, dtype=torch.long))


def test_sequential_tuple_semantics():
    class A(torch.nn.Module):
        def forward(self, x, y=None):
            return x + 1, y

    class B(torch.nn.Module):
        def forward(self, x, y):
            return (x, y)

    class C(torch.nn.Module):
        def forward(self, x, y):
            return x if y is None else x * y
    s = perceiver.Sequential(A(), B(), C())
    assert s(torch.tensor(1.0)).item() == 2.0
    assert s(torch.tensor(1.0), torch.tensor(3.0)).item() == 6.0
