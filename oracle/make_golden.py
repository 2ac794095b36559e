"""Generate tests/golden/*.npz from the UNMODIFIED reference (run in the build container only).

TEST INFRASTRUCTURE.  Usage:  python oracle/make_golden.py
Each fixture holds: the seeded inputs, the reference model's full state_dict (``sd/<key>``),
its forward output, a scalar loss and the gradient of that loss w.r.t. every parameter
(``grad/<key>``), all fp32 from torch CPU.  Configs are deliberately tiny so the fixtures stay
small; full-size configs are checked on the GPU box against oracle/perceiver_oracle.py, which
these fixtures pin.
"""
import json
import os
import sys

import numpy as np
import torch
import torch.nn.functional as F

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from oracle import reference_loader  # noqa: E402

OUT = os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "tests", "golden")


def _save(name, meta, arrays):
    arrays = {k: (v.detach().cpu().numpy() if isinstance(v, torch.Tensor) else np.asarray(v)) for k, v in arrays.items()}
    arrays["meta"] = np.frombuffer(json.dumps(meta).encode(), dtype=np.uint8)
    np.savez_compressed(os.path.join(OUT, name + ".npz"), **arrays)
    print(name, sum(a.nbytes for a in arrays.values()) // 1024, "KiB raw")


def _pack(model, arrays, loss):
    loss.backward()
    for k, v in model.state_dict().items():
        arrays["sd/" + k] = v
    for k, p in model.named_parameters():
        arrays["grad/" + k] = p.grad if p.grad is not None else torch.zeros_like(p)


def image_case(M, A, name, image_shape, bands, latent, layers, sa, heads, classes, batch, seed):
    torch.manual_seed(seed)
    enc = M.PerceiverEncoder(A.ImageInputAdapter(image_shape, bands), latent, layers,
                             num_cross_attention_heads=heads[0], num_self_attention_heads=heads[1],
                             num_self_attention_layers_per_block=sa)
    dec = M.PerceiverDecoder(A.ClassificationOutputAdapter(classes, num_output_channels=latent[1]), latent,
                             num_cross_attention_heads=heads[2])
    model = M.PerceiverIO(enc, dec)
    _perturb(model)
    x = torch.rand(batch, *image_shape) * 2 - 1
    y = torch.randint(0, classes, (batch,))
    logits = model(x)
    loss = F.cross_entropy(logits, y)
    arrays = {"x": x, "y": y, "logits": logits, "loss": loss}
    _pack(model, arrays, loss)
    meta = dict(kind="image", image_shape=list(image_shape), bands=bands, latent=list(latent), layers=layers,
                sa=sa, heads=list(heads), classes=classes)
    _save(name, meta, arrays)


def _perturb(model):
    # torch zero-inits every MHA bias and LayerNorm is (1,0): perturb so bias/affine paths are exercised
    g = torch.Generator().manual_seed(1234)
    with torch.no_grad():
        for k, p in model.named_parameters():
            if k.endswith("bias") or "norm" in k or k.endswith("module.0.weight"):
                p.add_(torch.randn(p.shape, generator=g) * 0.1)


def text_case(M, A, name, vocab, seq, C, latent, layers, sa, heads, batch, seed, clf_classes=None, l=None):
    torch.manual_seed(seed)
    l = l or seq
    enc = M.PerceiverEncoder(A.TextInputAdapter(vocab, seq, C), latent, layers,
                             num_cross_attention_heads=heads[0], num_self_attention_heads=heads[1],
                             num_self_attention_layers_per_block=sa)
    if clf_classes is None:
        out = A.TextOutputAdapter(vocab, seq, num_output_channels=latent[1])
    else:
        out = A.ClassificationOutputAdapter(clf_classes, num_output_channels=latent[1])
    dec = M.PerceiverDecoder(out, latent, num_cross_attention_heads=heads[2])
    model = M.PerceiverIO(enc, dec)
    _perturb(model)
    ids = torch.randint(3, vocab, (batch, l))
    lens = torch.randint(max(1, l // 3), l + 1, (batch,))
    lens[0] = l
    pad = torch.arange(l)[None, :] >= lens[:, None]
    ids[pad] = 0
    ids[torch.rand(batch, l) < 0.05] = 1
    ids[pad] = 0
    arrays = {"ids": ids, "pad": pad}
    if clf_classes is None:
        masking = M.TextMasking(vocab, 1, 2, 3)
        xm, labels = masking(ids.clone(), pad)
        labels[0, 0] = ids[0, 0]  # at least one supervised position
        logits = model(xm, pad)[:, :l, :]
        loss = F.cross_entropy(logits.transpose(1, 2), labels)  # 'b m c -> b c m', lightning.py:225-226
        arrays.update(ids_masked=xm, labels=labels)
    else:
        y = torch.randint(0, clf_classes, (batch,))
        logits = model(ids, pad)
        loss = F.cross_entropy(logits, y)
        arrays.update(y=y)
    arrays.update(logits=logits, loss=loss)
    _pack(model, arrays, loss)
    meta = dict(kind="text", vocab=vocab, seq=seq, C=C, latent=list(latent), layers=layers, sa=sa,
                heads=list(heads), clf_classes=clf_classes)
    _save(name, meta, arrays)


def masking_case(M, name, vocab, shape, seed):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, vocab, shape, generator=g)
    lens = torch.randint(shape[1] // 4, shape[1] + 1, (shape[0],), generator=g)
    pad = torch.arange(shape[1])[None, :] >= lens[:, None]
    ids[torch.rand(shape, generator=g) < 0.03] = 1
    ids[pad] = 0
    torch.manual_seed(seed + 1)
    state = torch.get_rng_state()
    xm, labels = M.TextMasking(vocab, 1, 2, 3)(ids.clone(), pad)
    _save(name, dict(kind="masking", vocab=vocab, seed=seed + 1),
          {"ids": ids, "pad": pad, "ids_masked": xm, "labels": labels, "rng_state": state})


def fourier_case(A, name, image_shape, bands):
    ad = A.ImageInputAdapter(image_shape, bands)
    torch.manual_seed(3)
    x = torch.rand(2, *image_shape)
    _save(name, dict(kind="fourier", image_shape=list(image_shape), bands=bands),
          {"position_encoding": ad.position_encoding, "x": x, "out": ad(x)})


def main():
    os.makedirs(OUT, exist_ok=True)
    M, A = reference_loader.load()
    torch.set_num_threads(1)
    image_case(M, A, "img_tiny", (6, 5, 1), 3, (8, 16), 2, 2, (2, 2, 1), 4, 3, seed=0)
    image_case(M, A, "img_rgb_1layer", (4, 4, 3), 2, (5, 8), 1, 1, (1, 2, 2), 3, 2, seed=1)
    text_case(M, A, "mlm_tiny", 50, 12, 16, (8, 16), 3, 2, (2, 2, 2), 4, seed=2)
    text_case(M, A, "mlm_short_l", 40, 16, 8, (4, 8), 2, 1, (1, 1, 1), 2, seed=5, l=9)
    text_case(M, A, "seqclf_tiny", 50, 12, 16, (8, 16), 2, 2, (4, 2, 1), 3, seed=3, clf_classes=2)
    masking_case(M, "masking_10003", 10003, (8, 96), seed=7)
    fourier_case(A, "fourier_6x5_f3", (6, 5, 1), 3)
    fourier_case(A, "fourier_3d", (3, 4, 2, 2), 2)
    # MNIST-sized table: keep a strided sample of rows only (fixture size)
    ad = A.ImageInputAdapter((28, 28, 1), 32)
    _save("fourier_28x28_f32_rows", dict(kind="fourier_rows", image_shape=[28, 28, 1], bands=32, stride=37),
          {"rows": ad.position_encoding[::37], "sum": ad.position_encoding.double().sum()})


if __name__ == "__main__":
    main()
