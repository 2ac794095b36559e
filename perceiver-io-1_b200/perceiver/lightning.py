"""Task wrappers with the hyper-parameter surface of the reference's perceiver/lightning.py
(LitImageClassifier :88-126, LitTextClassifier :129-171, LitMaskedLanguageModel :174-256), without the
pytorch_lightning / torchmetrics dependency: plain nn.Modules that build the same models from the same
keyword arguments and expose the same ``forward(batch)`` / ``step(batch)`` / ``configure_optimizers()``.
A Lightning ``Trainer`` is not needed to run them (see bench.py for a hand-written step) and is out of
scope here; the three wiring bugs of the fork are repaired behind the same API:
``TextMasking`` gets its token ids, ``PerceiverMLM`` unpacks the encoder tuple, and the text classifier
reaches its encoder through the ``PerceiverIO`` Sequential.
"""
import importlib
from types import SimpleNamespace
from typing import Any, List, Optional, Tuple

import torch
import torch.nn as nn

from . import functional as F
from .adapter import ClassificationOutputAdapter, ImageInputAdapter, TextInputAdapter, TextOutputAdapter
from .model import PerceiverDecoder, PerceiverEncoder, PerceiverIO, PerceiverMLM, TextMasking
from .tokenizer import MASK_TOKEN_ID, SPECIAL_TOKENS, UNK_TOKEN_ID
from .utils import freeze


def instantiate_class(args, init: dict):
    """``{'class_path': 'torch.optim.AdamW', 'init_args': {...}}`` -> object (what LightningCLI's helper does)."""
    module, _, name = init['class_path'].rpartition('.')
    cls = getattr(importlib.import_module(module), name)
    args = args if isinstance(args, tuple) else (args,)
    return cls(*args, **init.get('init_args', {}))


class LitModel(nn.Module):
    def __init__(self,
                 optimizer_init: Optional[dict] = None,
                 scheduler_init: Optional[dict] = None,
                 num_latents: int = 64,
                 num_latent_channels: int = 64,
                 num_encoder_layers: int = 3,
                 num_encoder_cross_attention_heads: int = 4,
                 num_encoder_self_attention_heads: int = 4,
                 num_encoder_self_attention_layers_per_block: int = 6,
                 num_decoder_cross_attention_heads: int = 4,
                 dropout: float = 0.0,
                 **task_hparams: Any):
        super().__init__()
        self.hparams = SimpleNamespace(
            optimizer_init=optimizer_init, scheduler_init=scheduler_init, num_latents=num_latents,
            num_latent_channels=num_latent_channels, num_encoder_layers=num_encoder_layers,
            num_encoder_cross_attention_heads=num_encoder_cross_attention_heads,
            num_encoder_self_attention_heads=num_encoder_self_attention_heads,
            num_encoder_self_attention_layers_per_block=num_encoder_self_attention_layers_per_block,
            num_decoder_cross_attention_heads=num_decoder_cross_attention_heads, dropout=dropout, **task_hparams)

    @property
    def latent_shape(self):
        return self.hparams.num_latents, self.hparams.num_latent_channels

    def _encoder(self, input_adapter):
        h = self.hparams
        return PerceiverEncoder(input_adapter=input_adapter, latent_shape=self.latent_shape, num_layers=h.num_encoder_layers,
                                num_cross_attention_heads=h.num_encoder_cross_attention_heads,
                                num_self_attention_heads=h.num_encoder_self_attention_heads,
                                num_self_attention_layers_per_block=h.num_encoder_self_attention_layers_per_block,
                                dropout=h.dropout)

    def _decoder(self, output_adapter):
        h = self.hparams
        return PerceiverDecoder(output_adapter=output_adapter, latent_shape=self.latent_shape,
                                num_cross_attention_heads=h.num_decoder_cross_attention_heads, dropout=h.dropout)

    def configure_optimizers(self):
        if self.hparams.optimizer_init is None:
            raise ValueError('optimizer_init was not given')
        optimizer = instantiate_class(self.parameters(), self.hparams.optimizer_init)
        if self.hparams.scheduler_init is None:
            return optimizer
        scheduler = instantiate_class(optimizer, self.hparams.scheduler_init)
        return {'optimizer': optimizer, 'lr_scheduler': {'scheduler': scheduler, 'interval': 'step', 'frequency': 1}}


class LitClassifier(LitModel):
    def step(self, batch):
        logits, y = self(batch)
        loss = F.cross_entropy(logits, y)
        acc = (logits.argmax(dim=-1) == y).float().mean()
        return loss, acc


class LitImageClassifier(LitClassifier):
    def __init__(self, image_shape: Tuple[int, int, int], num_classes: int, *args: Any, num_frequency_bands: int = 32,
                 **kwargs: Any):
        super().__init__(*args, image_shape=tuple(image_shape), num_classes=num_classes,
                         num_frequency_bands=num_frequency_bands, **kwargs)
        h = self.hparams
        encoder = self._encoder(ImageInputAdapter(image_shape=h.image_shape, num_frequency_bands=h.num_frequency_bands))
        decoder = self._decoder(ClassificationOutputAdapter(num_classes=h.num_classes,
                                                            num_output_channels=h.num_latent_channels))
        self.model = PerceiverIO(encoder, decoder)

    def forward(self, batch):
        x, y = batch
        return self.model(x), y


class LitTextClassifier(LitClassifier):
    def __init__(self, num_classes: int, vocab_size: int, max_seq_len: int, *args: Any, freeze_encoder: bool = False,
                 mlm_ckpt: Optional[str] = None, clf_ckpt: Optional[str] = None, **kwargs: Any):
        super().__init__(*args, num_classes=num_classes, vocab_size=vocab_size, max_seq_len=max_seq_len,
                         freeze_encoder=freeze_encoder, mlm_ckpt=mlm_ckpt, clf_ckpt=clf_ckpt, **kwargs)
        h = self.hparams
        encoder = self._encoder(TextInputAdapter(vocab_size=h.vocab_size, max_seq_len=h.max_seq_len,
                                                 num_input_channels=h.num_latent_channels))
        decoder = self._decoder(ClassificationOutputAdapter(num_classes=h.num_classes,
                                                            num_output_channels=h.num_latent_channels))
        self.model = PerceiverIO(encoder, decoder)
        if mlm_ckpt is not None:  # warm start from a masked-LM checkpoint: its 'model.encoder.*' weights
            self.encoder.load_state_dict(_sub_state_dict(_load_ckpt(mlm_ckpt), 'model.encoder.'))
        elif clf_ckpt is not None:
            self.model.load_state_dict(_sub_state_dict(_load_ckpt(clf_ckpt), 'model.'))
        if freeze_encoder:
            freeze(self.encoder)

    @property
    def encoder(self):
        return self.model[0]  # PerceiverIO is Sequential(encoder, decoder)

    def forward(self, batch):
        y, x, x_mask = batch
        return self.model(x, x_mask), y


class LitMaskedLanguageModel(LitModel):
    def __init__(self, vocab_size: int, max_seq_len: int, *args: Any, masked_samples: Optional[List[str]] = None,
                 num_predictions: int = 3, **kwargs: Any):
        super().__init__(*args, vocab_size=vocab_size, max_seq_len=max_seq_len, masked_samples=masked_samples,
                         num_predictions=num_predictions, **kwargs)
        h = self.hparams
        encoder = self._encoder(TextInputAdapter(vocab_size=h.vocab_size, max_seq_len=h.max_seq_len,
                                                 num_input_channels=h.num_latent_channels))
        decoder = self._decoder(TextOutputAdapter(vocab_size=h.vocab_size, max_seq_len=h.max_seq_len,
                                                  num_output_channels=h.num_latent_channels))
        masking = TextMasking(vocab_size=h.vocab_size, unk_token_id=UNK_TOKEN_ID, mask_token_id=MASK_TOKEN_ID,
                              num_special_tokens=len(SPECIAL_TOKENS))
        self.model = PerceiverMLM(encoder, decoder, masking)

    def forward(self, batch):
        _, x, x_mask = batch
        return self.model(x, x_mask)

    def step(self, batch):
        """Masked-LM loss of a (y, ids, pad_mask) batch; in bf16 mode through the fused vocabulary-projection
        + cross-entropy operator (same value and gradients, logits never stored)."""
        _, x, x_mask = batch
        if F.get_compute_mode() == 1 and torch.is_grad_enabled():
            loss, _ = self.model(x, x_mask, return_loss=True)
            return loss
        logits, labels = self.model(x, x_mask)
        return F.cross_entropy(logits, labels)


def _load_ckpt(path):
    ckpt = torch.load(path, map_location='cpu')
    return ckpt.get('state_dict', ckpt)


def _sub_state_dict(sd, prefix):
    return {k[len(prefix):]: v for k, v in sd.items() if k.startswith(prefix)}
