"""The task wrappers (reference perceiver/lightning.py: LitImageClassifier :88-126, LitTextClassifier :129-171,
LitMaskedLanguageModel :174-256) driven on the GPU: `step()` of each against the oracle on the same weights and batch,
`configure_optimizers()` + a few optimizer steps, and `predict_masked_samples` (reference utils.py:22-43)."""
import pytest
import torch

from gpu_util import BF16_TOL, FP32_TOL, assert_rel
from oracle import perceiver_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
OPT = {"class_path": "torch.optim.AdamW", "init_args": {"lr": 1e-3}}


@pytest.fixture(params=["fp32", "bf16"])
def mode(request):
    import perceiver
    perceiver.set_compute_mode(request.param)
    yield request.param
    perceiver.set_compute_mode("fp32")


def _sd(model):
    return {k: v.detach().cpu().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}


def test_lit_image_classifier_step(mode):
    from perceiver import LitImageClassifier
    torch.manual_seed(0)
    lit = LitImageClassifier(image_shape=(12, 12, 1), num_classes=10, num_frequency_bands=8, num_latents=16, num_latent_channels=64,
                             num_encoder_layers=2, num_encoder_self_attention_layers_per_block=2, num_decoder_cross_attention_heads=1,
                             optimizer_init=OPT).to(DEV)
    x = torch.rand(6, 12, 12, 1) * 2 - 1
    y = torch.randint(0, 10, (6,))
    sd = _sd(lit.model)
    ref_logits = O.perceiver_io(sd, x, None, kind="image", num_layers=2, cross_heads=4, self_heads=4, sa_layers=2, dec_heads=1)
    ref_loss = O.cross_entropy(ref_logits, y)
    loss, acc = lit.step((x.to(DEV), y.to(DEV)))
    tol = FP32_TOL if mode == "fp32" else BF16_TOL
    assert_rel(loss, ref_loss, tol, "loss")
    assert abs(acc.item() - (ref_logits.argmax(-1) == y).float().mean().item()) <= (0.0 if mode == "fp32" else 0.34)
    opt = lit.configure_optimizers()
    first = loss.item()
    for _ in range(8):  # the wrapper trains: the loss of a fixed batch falls
        opt.zero_grad(set_to_none=True)
        loss, _ = lit.step((x.to(DEV), y.to(DEV)))
        loss.backward()
        opt.step()
    assert loss.item() < first


def test_lit_text_classifier_step_and_frozen_encoder(mode):
    from perceiver import LitTextClassifier
    torch.manual_seed(1)
    kw = dict(num_classes=2, vocab_size=503, max_seq_len=64, num_latents=16, num_latent_channels=64, num_encoder_layers=2,
              num_encoder_self_attention_layers_per_block=2, num_decoder_cross_attention_heads=1, optimizer_init=OPT)
    lit = LitTextClassifier(**kw).to(DEV)
    ids = torch.randint(3, 503, (5, 64))
    pad = torch.arange(64)[None, :] >= torch.tensor([64, 50, 33, 64, 20])[:, None]
    ids[pad] = 0
    y = torch.randint(0, 2, (5,))
    sd = _sd(lit.model)
    ref_loss = O.cross_entropy(O.perceiver_io(sd, ids, pad, kind="text", num_layers=2, cross_heads=4, self_heads=4, sa_layers=2,
                                              dec_heads=1), y)
    loss, _ = lit.step((y.to(DEV), ids.to(DEV), pad.to(DEV)))
    assert_rel(loss, ref_loss, FP32_TOL if mode == "fp32" else BF16_TOL, "loss")
    loss.backward()
    assert all(p.grad is not None for p in lit.parameters())
    frozen = LitTextClassifier(freeze_encoder=True, **kw).to(DEV)
    loss, _ = frozen.step((y.to(DEV), ids.to(DEV), pad.to(DEV)))
    loss.backward()
    assert all(p.grad is None for p in frozen.encoder.parameters())
    assert all(p.grad is not None for p in frozen.model[1].parameters())


def test_lit_masked_language_model_step(mode):
    from perceiver import LitMaskedLanguageModel
    torch.manual_seed(2)
    lit = LitMaskedLanguageModel(vocab_size=503, max_seq_len=64, num_latents=16, num_latent_channels=64, num_encoder_layers=2,
                                 num_encoder_self_attention_layers_per_block=2, num_decoder_cross_attention_heads=2,
                                 optimizer_init=OPT,
                                 scheduler_init={"class_path": "torch.optim.lr_scheduler.OneCycleLR",
                                                 "init_args": {"max_lr": 1e-3, "total_steps": 20}}).to(DEV)
    ids = torch.randint(3, 503, (6, 64), device=DEV)
    pad = torch.arange(64, device=DEV)[None, :] >= torch.tensor([64, 50, 33, 64, 20, 64], device=DEV)[:, None]
    ids[pad] = 0
    # step() masks in place with torch's generator: replay the same draws for the oracle
    state = torch.cuda.get_rng_state()
    loss = lit.step((None, ids.clone(), pad))
    torch.cuda.set_rng_state(state)
    masked, labels = lit.model.masking(ids.clone(), pad)
    sd = _sd(lit.model)
    ref_logits = O.mlm_forward(sd, masked.cpu(), pad.cpu(), num_layers=2, cross_heads=4, self_heads=4, sa_layers=2, dec_heads=2)
    ref_loss = O.cross_entropy(ref_logits, labels.cpu())
    assert_rel(loss, ref_loss, FP32_TOL if mode == "fp32" else BF16_TOL, "loss")
    cfg = lit.configure_optimizers()
    opt, sched = cfg["optimizer"], cfg["lr_scheduler"]["scheduler"]
    first = None
    for _ in range(6):
        opt.zero_grad(set_to_none=True)
        torch.manual_seed(3)  # the same masking every step: the loss of the fixed batch falls
        loss = lit.step((None, ids.clone(), pad))
        first = loss.item() if first is None else first
        loss.backward()
        opt.step()
        sched.step()
    assert loss.item() < first


def test_predict_masked_samples_runs_on_the_gpu_model():
    """utils.predict_masked_samples (reference utils.py:22-43): top-k fill-ins of the [MASK] positions"""
    from perceiver import LitMaskedLanguageModel, predict_masked_samples
    from perceiver.tokenizer import MASK_TOKEN_ID, PAD_TOKEN_ID

    class Tok:  # the slice of the HuggingFace tokenizer API the function uses
        def enable_padding(self, **kw):
            self.pad = kw

        def no_padding(self):
            self.pad = None

        def encode_batch(self, samples):
            import types
            rows = [[3 + (hash(w) % 400) if w != "[MASK]" else MASK_TOKEN_ID for w in s.split()] for s in samples]
            n = max(len(r) for r in rows)
            return [types.SimpleNamespace(ids=r + [PAD_TOKEN_ID] * (n - len(r))) for r in rows]

        def decode(self, row, skip_special_tokens=True):
            return " ".join(str(t) for t in row if t > 2)

        def token_to_id(self, t):
            return {"[PAD]": PAD_TOKEN_ID, "[MASK]": MASK_TOKEN_ID}[t]

    torch.manual_seed(4)
    lit = LitMaskedLanguageModel(vocab_size=503, max_seq_len=32, num_latents=8, num_latent_channels=64, num_encoder_layers=1,
                                 num_encoder_self_attention_layers_per_block=1).to(DEV).eval()
    samples = ["a b [MASK] d e", "f [MASK] h"]
    out = predict_masked_samples(samples, lambda s: _encode(Tok(), s), Tok(), lit.model, device=DEV, num_predictions=3)
    assert len(out) == 2 and all(len(o) == 3 for o in out)


def _encode(tok, samples):
    enc = tok.encode_batch(samples)
    ids = torch.tensor([e.ids for e in enc])
    return ids, ids == 0
