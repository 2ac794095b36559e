"""Host logic of the Lightning-free task wrappers: hparams -> model mapping as in the reference's
lightning.py, script defaults (scripts/img_clf.py:14-22, mlm.py:23-25, seq_clf.py:15-22), optimizer
construction from {class_path, init_args}.  No GPU."""
import torch

from perceiver import LitImageClassifier, LitMaskedLanguageModel, LitTextClassifier, PerceiverIO, PerceiverMLM


def test_image_classifier_config1_defaults():
    m = LitImageClassifier(image_shape=(28, 28, 1), num_classes=10, num_latents=32, num_latent_channels=128,
                           num_encoder_layers=3, num_encoder_self_attention_layers_per_block=3,
                           num_decoder_cross_attention_heads=1)
    assert isinstance(m.model, PerceiverIO)
    enc, dec = m.model[0], m.model[1]
    assert enc.input_adapter.num_input_channels == 131 and enc.latent.shape == (32, 128)
    assert len(enc.layer_1[1]) == 3 and enc.num_layers == 3
    assert dec.output.shape == (1, 128) and dec.output_adapter.linear.weight.shape == (10, 128)
    sd = m.state_dict()
    assert 'model.0.layer_n.1.2.1.module.3.weight' in sd and 'model.0.input_adapter.position_encoding' in sd


def test_mlm_config2_parameter_count_matches_survey_probe():
    m = LitMaskedLanguageModel(vocab_size=10003, max_seq_len=512, num_latents=64, num_latent_channels=256)
    assert isinstance(m.model, PerceiverMLM)
    assert m.model.masking.unk_token_id == 1 and m.model.masking.mask_token_id == 2 and m.model.masking.num_special_tokens == 3
    n = sum(p.numel() for p in m.parameters())
    assert n == 11_348_243  # gradient set of config 2


def test_text_classifier_reaches_encoder_and_freezes_it(tmp_path):
    mlm = LitMaskedLanguageModel(vocab_size=50, max_seq_len=16, num_latents=8, num_latent_channels=16,
                                 num_encoder_self_attention_layers_per_block=1, num_encoder_layers=2)
    ckpt = tmp_path / "mlm.ckpt"
    torch.save({"state_dict": mlm.state_dict()}, ckpt)
    clf = LitTextClassifier(num_classes=2, vocab_size=50, max_seq_len=16, num_latents=8, num_latent_channels=16,
                            num_encoder_self_attention_layers_per_block=1, num_encoder_layers=2,
                            num_decoder_cross_attention_heads=1, freeze_encoder=True, mlm_ckpt=str(ckpt))
    for k, v in mlm.model.encoder.state_dict().items():
        assert torch.equal(clf.encoder.state_dict()[k], v), k
    assert all(not p.requires_grad for p in clf.encoder.parameters())
    assert all(p.requires_grad for p in clf.model[1].parameters())


def test_configure_optimizers_from_class_path():
    m = LitImageClassifier(image_shape=(6, 5, 1), num_classes=3, num_latents=4, num_latent_channels=8,
                           num_encoder_layers=1, num_encoder_cross_attention_heads=1, num_encoder_self_attention_heads=1,
                           num_encoder_self_attention_layers_per_block=1, num_decoder_cross_attention_heads=1,
                           num_frequency_bands=2,
                           optimizer_init={'class_path': 'torch.optim.AdamW', 'init_args': {'lr': 1e-3, 'weight_decay': 0.01}},
                           scheduler_init={'class_path': 'torch.optim.lr_scheduler.OneCycleLR',
                                           'init_args': {'max_lr': 1e-3, 'total_steps': 100, 'pct_start': 0.1}})
    cfg = m.configure_optimizers()
    assert isinstance(cfg['optimizer'], torch.optim.AdamW)
    assert cfg['lr_scheduler']['interval'] == 'step'
