"""Perceiver IO encoder / decoder — drop-in for the reference's perceiver/model.py.

Same class names, constructor signatures, forward conventions, exceptions and state_dict keys;
every forward runs the fused sm_100a operators of libpio_b200.so (functional.py).  Differences
that are deliberate and documented in DESIGN.md:
  * the key/value projection of the shared ``layer_n`` is computed once per forward and reused
    by its (num_layers-1) applications (same weights, same input — reference model.py:186-187);
  * ``PerceiverMLM.forward`` uses the wiring the reference evidently intends (its own raises a
    TypeError) and only decodes the ``l`` queries it returns;
  * dropout > 0 (training mode) uses the kernels' own counter-based masks — same sites and scaling as
    the reference (attention probabilities, Residual branch output), seeded from torch.initial_seed(),
    but not torch's Philox stream, so parity at p > 0 is statistical; ``attn_mask`` is rejected (no
    caller of the reference ever passes one).
"""
from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import functional as F
from .adapter import InputAdapter, OutputAdapter
from .layers import LayerNorm, Linear, MultiheadAttentionParams
from .tokenizer import MASK_TOKEN, SPECIAL_TOKENS, UNK_TOKEN
from .utils import Sequential


def _scoped(forward):
    """run a model-level forward inside a bf16 weight-shadow scope (functional.shadow_scope)"""
    import functools

    @functools.wraps(forward)
    def wrapped(self, *args, **kwargs):
        with F.shadow_scope(self):
            return forward(self, *args, **kwargs)
    return wrapped


def _active_p(module: nn.Module, p: float) -> float:
    """dropout probability in effect: the module's p in training mode, 0 in eval mode"""
    return float(p) if (module.training and p > 0.0) else 0.0


def _no_attn_mask(attn_mask):
    if attn_mask is not None:
        raise NotImplementedError('attn_mask is not supported (no caller of the reference passes one)')


class MLP(Sequential):
    """LN -> Linear -> GELU(erf) -> Linear with the reference's child indices 0,1,(2),3."""

    def __init__(self, num_channels: int):
        super().__init__(LayerNorm(num_channels), Linear(num_channels, num_channels), nn.GELU(),
                         Linear(num_channels, num_channels))

    def _params(self):
        return (self[0].weight, self[0].bias, self[1].weight, self[1].bias, self[3].weight, self[3].bias)

    def forward(self, x, residual: bool = False, dropout_p: float = 0.0):
        """``dropout_p`` is the enclosing Residual's dropout (only applied together with ``residual``)."""
        seed = F.next_dropout_seed() if dropout_p else 0
        return F.MlpFn.apply(x, *self._params(), residual, dropout_p, seed)


def mlp(num_channels: int):
    return MLP(num_channels)


class MultiHeadAttention(nn.Module):
    def __init__(self, num_q_channels: int, num_kv_channels: int, num_heads: int, dropout: float):
        super().__init__()
        self.attention = MultiheadAttentionParams(embed_dim=num_q_channels, num_heads=num_heads,
                                                  kdim=num_kv_channels, vdim=num_kv_channels, dropout=dropout)


class CrossAttention(nn.Module):
    def __init__(self, num_q_channels: int, num_kv_channels: int, num_heads: int, dropout: float):
        super().__init__()
        self.q_norm = LayerNorm(num_q_channels)
        self.kv_norm = LayerNorm(num_kv_channels)
        self.attention = MultiHeadAttention(num_q_channels=num_q_channels, num_kv_channels=num_kv_channels,
                                            num_heads=num_heads, dropout=dropout)
        self.dropout_p = dropout

    def project_kv(self, x_kv):
        """kv_norm + K/V in-projection -> (B, M, 2E); reusable across applications of this layer."""
        a = self.attention.attention
        w_packed, _, w_k, w_v = a.packed()
        return F.KVProjectFn.apply(x_kv, self.kv_norm.weight, self.kv_norm.bias, a.in_proj_bias, w_packed, w_k, w_v,
                                   a.embed_dim)

    def forward(self, x_q, x_kv, pad_mask=None, attn_mask=None, kv=None, residual: bool = False, dropout_p=None):
        _no_attn_mask(attn_mask)
        p = _active_p(self, self.dropout_p) if dropout_p is None else dropout_p
        a = self.attention.attention
        if kv is None:
            kv = self.project_kv(x_kv)
        w_packed, w_q, _, _ = a.packed()
        seed = F.next_dropout_seed() if p else 0
        deferred = residual == 2  # (the consumer applies the out-projection: see SelfAttentionBlock.forward)
        return F.CrossAttendFn.apply(x_q, kv, pad_mask, self.q_norm.weight, self.q_norm.bias, a.in_proj_bias, w_packed,
                                     w_q, None if deferred else a.out_proj.weight, None if deferred else a.out_proj.bias,
                                     a.num_heads, residual, p, seed)


class SelfAttention(nn.Module):
    def __init__(self, num_channels: int, num_heads: int, dropout: float):
        super().__init__()
        self.norm = LayerNorm(num_channels)
        self.attention = MultiHeadAttention(num_q_channels=num_channels, num_kv_channels=num_channels,
                                            num_heads=num_heads, dropout=dropout)
        self.dropout_p = dropout

    def _params(self):
        a = self.attention.attention
        return (self.norm.weight, self.norm.bias, a.in_proj_weight, a.in_proj_bias, a.out_proj.weight, a.out_proj.bias)

    def forward(self, x, pad_mask=None, attn_mask=None, residual: bool = False, dropout_p=None):
        _no_attn_mask(attn_mask)
        p = _active_p(self, self.dropout_p) if dropout_p is None else dropout_p
        if pad_mask is not None:
            raise NotImplementedError('SelfAttention with a pad mask is never reached in the reference model')
        seed = F.next_dropout_seed() if p else 0
        return F.SelfAttnFn.apply(x, *self._params(), self.attention.attention.num_heads, residual, p, seed)


class Residual(nn.Module):
    """``dropout(module(*args)) + args[0]``; for the three module kinds of the model the add is
    fused into the operator's last GEMM epilogue."""

    def __init__(self, module: nn.Module, dropout: float):
        super().__init__()
        self.module = module
        self.dropout = nn.Dropout(p=dropout)
        self.dropout_p = dropout

    def forward(self, *args, **kwargs):
        if isinstance(self.module, (CrossAttention, SelfAttention, MLP)):
            # one p for the module's own attention dropout and for this Residual's dropout, like the
            # reference's cross_attention_layer / self_attention_layer constructors (model.py:29-40)
            return self.module(*args, residual=True, dropout_p=_active_p(self, self.dropout_p), **kwargs)
        return self.dropout(self.module(*args, **kwargs)) + args[0]


def cross_attention_layer(num_q_channels: int, num_kv_channels: int, num_heads: int, dropout: float):
    return Sequential(Residual(CrossAttention(num_q_channels, num_kv_channels, num_heads, dropout), dropout),
                      Residual(mlp(num_q_channels), dropout))


def self_attention_layer(num_channels: int, num_heads: int, dropout: float):
    return Sequential(Residual(SelfAttention(num_channels, num_heads, dropout), dropout),
                      Residual(mlp(num_channels), dropout))


class SelfAttentionBlock(Sequential):
    """S self-attention layers; forward hands all of them to one fused call.  ``pre``: the Residual(mlp) of the
    cross-attention layer in front of the block (reference model.py:29-33), run by the same call as an MLP-only first
    layer when the fused cluster kernels take the block (see `takes_pre_mlp`)."""

    def takes_pre_mlp(self, x, pre):
        import os
        if F.get_compute_mode() != F.N.MODE_BF16 or os.environ.get("PIO_NO_PRE_MLP") is not None or not x.is_cuda:
            return False
        if _active_p(self, self[0][0].dropout_p) or _active_p(pre, pre.dropout_p):
            return False
        b, n, c = x.shape
        heads = self[0][0].module.attention.attention.num_heads
        return bool(F.N.lib().pio_latent_block_cluster_ok(len(self) + 1, int(b), int(n), int(c), int(heads)))

    def forward(self, x, pre=None, pre_attn=None, o16=None):
        """``pre_attn`` / ``o16``: the CrossAttention in front and its attention output before the out-projection
        (CrossAttendFn with residual=2): the first layer then also applies that out-projection and its residual."""
        params = []
        if pre is not None:
            if pre_attn is not None:
                a = pre_attn.attention.attention
                params.extend([None, None, None, None, a.out_proj.weight, a.out_proj.bias])
            else:
                params.extend([None] * 6)
            params.extend(pre.module._params())
        for layer in self:
            sa, ff = layer[0], layer[1]
            params.extend(sa.module._params())
            params.extend(ff.module._params())
        heads = self[0][0].module.attention.attention.num_heads
        p = _active_p(self, self[0][0].dropout_p)
        seed = F.next_dropout_seed() if p else 0
        return F.LatentBlockFn.apply(x, o16, heads, p, seed, *params)


def self_attention_block(num_layers: int, num_channels: int, num_heads: int, dropout: float):
    return SelfAttentionBlock(*[self_attention_layer(num_channels, num_heads, dropout) for _ in range(num_layers)])


class PerceiverEncoder(nn.Module):
    def __init__(self,
                 input_adapter: InputAdapter,
                 latent_shape: Tuple[int, int],
                 num_layers: int,
                 num_cross_attention_heads: int = 4,
                 num_self_attention_heads: int = 4,
                 num_self_attention_layers_per_block: int = 2,
                 dropout: float = 0.0):
        """Generic Perceiver IO encoder: input adapter -> (B, M, C_in); latent array (N, C) cross-attends
        to it once per layer, each followed by a block of latent self-attention layers.  Layers 2..L
        share one set of weights (``layer_n``)."""
        super().__init__()
        self.input_adapter = input_adapter
        self.num_layers = num_layers
        c = latent_shape[1]

        def perceiver_layer():
            return Sequential(
                cross_attention_layer(c, input_adapter.num_input_channels, num_cross_attention_heads, dropout),
                self_attention_block(num_self_attention_layers_per_block, c, num_self_attention_heads, dropout))

        self.layer_1 = perceiver_layer()
        if num_layers > 1:
            self.layer_n = perceiver_layer()
        self.latent = nn.Parameter(torch.empty(*latent_shape))
        with torch.no_grad():
            self.latent.normal_(0.0, 0.02).clamp_(-2.0, 2.0)

    @staticmethod
    def _apply_layer(layer, x_latent, x, pad_mask, attn_mask, kv=None):
        cross, block = layer[0], layer[1]
        if PerceiverEncoder._defer_out_proj(cross, block, x_latent):
            ca = cross[0].module
            o16 = ca(x_latent, x, pad_mask, attn_mask, kv=kv, residual=2, dropout_p=0.0)  # attention output, no out-projection
            return block(x_latent, pre=cross[1], pre_attn=ca, o16=o16)
        x_latent = cross[0](x_latent, x, pad_mask, attn_mask, kv=kv)
        if isinstance(block, SelfAttentionBlock) and block.takes_pre_mlp(x_latent, cross[1]):
            return block(x_latent, pre=cross[1])  # the cross-attention layer's MLP rides on the block's fused kernels
        x_latent = cross[1](x_latent)
        return block(x_latent)

    @staticmethod
    def _defer_out_proj(cross, block, x_latent):
        """The cross-attention layer's out-projection + residual can ride on the block's fused kernels too (tensor-core
        attention head sizes, no dropout, the block takes the layer's MLP)."""
        import os
        ca = cross[0].module
        a = ca.attention.attention
        return (isinstance(block, SelfAttentionBlock) and os.environ.get("PIO_NO_PRE_PROJ") is None and
                not _active_p(cross[0], cross[0].dropout_p) and (a.embed_dim // a.num_heads) in (32, 64, 128) and
                block.takes_pre_mlp(x_latent, cross[1]))

    @_scoped
    def forward(self, x, pad_mask=None, attn_mask=None):
        b = x.shape[0]
        x = self.input_adapter(x)
        x_latent = F.BroadcastRowsFn.apply(self.latent, b, self.latent.shape[0])
        x_latent = self._apply_layer(self.layer_1, x_latent, x, pad_mask, attn_mask)
        if self.num_layers > 1:
            kv_n = self.layer_n[0][0].module.project_kv(x)  # once for all repeats of the shared layer
            for _ in range(self.num_layers - 1):
                x_latent = self._apply_layer(self.layer_n, x_latent, x, pad_mask, attn_mask, kv=kv_n)
        return x_latent, pad_mask


class PerceiverDecoder(nn.Module):
    def __init__(self,
                 output_adapter: OutputAdapter,
                 latent_shape: Tuple[int, int],
                 num_cross_attention_heads: int = 4,
                 dropout: float = 0.0):
        """Generic Perceiver IO decoder: a learned (K, C_out) query array cross-attends to the latents;
        the output adapter maps (B, K, C_out) to the task output."""
        super().__init__()
        c_out = output_adapter.output_shape[-1]
        self.output_adapter = output_adapter
        self.latent_shape = latent_shape
        self.cross_attention = cross_attention_layer(num_q_channels=c_out, num_kv_channels=latent_shape[1],
                                                     num_heads=num_cross_attention_heads, dropout=dropout)
        self.output = nn.Parameter(torch.empty(*output_adapter.output_shape))
        with torch.no_grad():
            self.output.normal_(0.0, 0.02).clamp_(-2.0, 2.0)

    def decode(self, x, num_queries: Optional[int] = None):
        """(B, K, C_out) decoder output before the output adapter."""
        b, *d = x.shape
        if tuple(d) != tuple(self.latent_shape):
            raise ValueError(f'Latent shape {tuple(d)} different from required shape {self.latent_shape}')
        q = F.BroadcastRowsFn.apply(self.output, b, self.output.shape[0] if num_queries is None else num_queries)
        return self.cross_attention(q, x)  # no mask on the latents, as in the reference

    @_scoped
    def forward(self, x, pad_mask=None, num_queries: Optional[int] = None):
        return self.output_adapter(self.decode(x, num_queries))  # pad_mask is accepted and unused, as in the reference


class TextMasking(nn.Module):
    def __init__(self, vocab_size: int, unk_token_id: int, mask_token_id: int, num_special_tokens: int,
                 mask_p: float = 0.15):
        """BERT-style 80/10/10 masking (https://arxiv.org/abs/1810.04805)."""
        super().__init__()
        self.vocab_size = vocab_size
        self.unk_token_id = unk_token_id
        self.mask_token_id = mask_token_id
        self.num_special_tokens = num_special_tokens
        self.mask_p = mask_p

    @staticmethod
    def create(tokenizer, **kwargs):
        return TextMasking(vocab_size=tokenizer.get_vocab_size(),
                           unk_token_id=tokenizer.token_to_id(UNK_TOKEN),
                           mask_token_id=tokenizer.token_to_id(MASK_TOKEN),
                           num_special_tokens=len(SPECIAL_TOKENS), **kwargs)

    def forward(self, x, pad_mask, generator=None):
        labels = F.text_mask(x, pad_mask, self.vocab_size, self.unk_token_id, self.mask_token_id,
                             self.num_special_tokens, self.mask_p, generator=generator)
        return x, labels


class PerceiverMLM(nn.Module):
    def __init__(self, encoder: PerceiverEncoder, decoder: PerceiverDecoder, masking: TextMasking):
        super().__init__()
        self.encoder = encoder
        self.decoder = decoder
        self.masking = masking

    @_scoped
    def forward(self, x_input, pad_mask=None, masking=True, return_loss=False, labels=None):
        """-> (logits[:, :l], labels) like the reference.  ``return_loss=True`` (training) returns
        (loss, labels) instead: the vocabulary projection and the cross-entropy of
        LitMaskedLanguageModel.step (reference lightning.py:223-226) run as one fused operator and the
        (B, l, vocab) logits are never stored.  With ``masking=False`` the already masked batch can bring
        its own ``labels`` (e.g. when the masking step runs outside a captured CUDA graph)."""
        _, l = x_input.shape
        if masking:
            x_masked, x_labels = self.masking(x_input, pad_mask)
        else:
            x_masked, x_labels = x_input, labels
        x_latent, _ = self.encoder(x_masked, pad_mask)
        if return_loss:
            if x_labels is None:
                raise ValueError('return_loss=True needs labels: masking=True, or masking=False with labels=...')
            lin = self.decoder.output_adapter.linear
            hidden_c = self._decode_labelled(x_latent, x_labels, l)
            if hidden_c is not None:  # decoder evaluated for the labelled (sample, position) pairs only
                hidden, labels_c = hidden_c
                return F.linear_cross_entropy(hidden, lin.weight, lin.bias, labels_c), x_labels
            hidden = self.decoder.decode(x_latent, num_queries=l)
            return F.linear_cross_entropy(hidden, lin.weight, lin.bias, x_labels), x_labels
        x_logits = self.decoder(x_latent, pad_mask, num_queries=l)
        return x_logits, x_labels


    def _decode_labelled(self, x_latent, labels, l):
        """The training loss only reads the decoder at the positions that carry a label (an ignored position adds exactly
        zero to the loss and to every gradient) and each decoder query is independent of the others (reference
        model.py:229-237), so in bf16 mode the decoder layer runs on those ~10-15 % of the (sample, position) pairs only
        (functional.DecodeRowsFn; compacted on the device, count never read by the host).  -> (rows, labels of the rows),
        or None when the layer's shape / mode / dropout is outside what that operator covers."""
        import os
        dec = self.decoder
        if not (torch.is_grad_enabled() and F.get_compute_mode() == F.N.MODE_BF16) or os.environ.get("PIO_NO_COMPACT_DECODE"):
            return None
        layer = dec.cross_attention
        res_att, res_mlp = layer[0], layer[1]
        if not (isinstance(res_att, Residual) and isinstance(res_att.module, CrossAttention) and isinstance(res_mlp.module, MLP)):
            return None
        if _active_p(res_att, res_att.dropout_p) > 0.0 or _active_p(res_mlp, res_mlp.dropout_p) > 0.0:
            return None
        b, *d = x_latent.shape
        if tuple(d) != tuple(dec.latent_shape):
            raise ValueError(f'Latent shape {tuple(d)} different from required shape {dec.latent_shape}')
        ca = res_att.module
        a = ca.attention.attention
        if not F.decode_rows_supported(d[0], a.embed_dim, a.num_heads) or dec.output.shape[0] < l:
            return None
        kv = ca.project_kv(x_latent)
        idx, labels_c, count = F.label_compact(labels, dec.output_adapter.linear.weight.shape[0])
        w_packed, w_q, _, _ = a.packed()
        hidden = F.DecodeRowsFn.apply(dec.output, kv, idx, count, l, ca.q_norm.weight, ca.q_norm.bias, a.in_proj_bias, w_packed, w_q,
                                      a.out_proj.weight, a.out_proj.bias, a.num_heads, *res_mlp.module._params())
        return hidden, labels_c


class PerceiverIO(Sequential):
    def __init__(self, encoder: PerceiverEncoder, decoder: PerceiverDecoder):
        super().__init__(encoder, decoder)

    @_scoped
    def forward(self, *inputs):
        return super().forward(*inputs)
