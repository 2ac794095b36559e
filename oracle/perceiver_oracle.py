"""CPU oracle: first-principles restatement of the reference's Perceiver IO hot path.

TEST INFRASTRUCTURE ONLY (see oracle/__init__.py).  Plain torch tensor arithmetic on a flat
``state_dict`` with the reference's key names; gradients come from torch autograd over this
restated math.  Device-agnostic (runs on whatever device its inputs live on) so the bit-exact
integer paths (TextMasking) can be checked on the same device/generator as the CUDA path.

Parity pinning: the reference has no tests/golden vectors of its own and its
attention arithmetic lives in a third-party dependency, ``torch.nn.MultiheadAttention``
(reference pins pytorch=1.10, environment.yml:8; the container has torch 2.11.0).  This oracle
is pinned by ``tests/golden/*.npz`` — outputs of the unmodified reference run in the build
container by ``oracle/make_golden.py`` (committed) — see tests/test_oracle_golden.py.

Each function cites the reference lines it restates (paths relative to /root/reference).
"""
import math
from typing import Dict, Optional, Tuple

import torch

StateDict = Dict[str, torch.Tensor]


# --------------------------------------------------------------------------------------
# primitives (third-party arithmetic the reference delegates to torch)
# --------------------------------------------------------------------------------------

def layer_norm(x: torch.Tensor, w: torch.Tensor, b: torch.Tensor, eps: float = 1e-5) -> torch.Tensor:
    """nn.LayerNorm over the last dim, biased variance (model.py:22,89,90,108)."""
    mu = x.mean(dim=-1, keepdim=True)
    xc = x - mu
    var = (xc * xc).mean(dim=-1, keepdim=True)
    return xc * torch.rsqrt(var + eps) * w + b


def gelu_erf(x: torch.Tensor) -> torch.Tensor:
    """nn.GELU() default = exact erf form (model.py:24)."""
    return 0.5 * x * (1.0 + torch.erf(x * (1.0 / math.sqrt(2.0))))


def linear(x: torch.Tensor, w: torch.Tensor, b: Optional[torch.Tensor]) -> torch.Tensor:
    y = x @ w.t()
    return y if b is None else y + b


def mha_weights(sd: StateDict, prefix: str):
    """Packed vs split in-projection as torch stores it (model.py:66-71): packed
    ``in_proj_weight (3E,E)`` when kdim == vdim == E, else ``{q,k,v}_proj_weight``; the bias is
    always packed ``in_proj_bias (3E,)``."""
    bias = sd[prefix + "in_proj_bias"]
    e = bias.shape[0] // 3
    if prefix + "in_proj_weight" in sd:
        w = sd[prefix + "in_proj_weight"]
        wq, wk, wv = w[:e], w[e:2 * e], w[2 * e:]
    else:
        wq, wk, wv = sd[prefix + "q_proj_weight"], sd[prefix + "k_proj_weight"], sd[prefix + "v_proj_weight"]
    return wq, wk, wv, bias[:e], bias[e:2 * e], bias[2 * e:], sd[prefix + "out_proj.weight"], sd[prefix + "out_proj.bias"]


def multi_head_attention(sd: StateDict, prefix: str, x_q: torch.Tensor, x_kv: torch.Tensor,
                         num_heads: int, pad_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``MultiHeadAttention.forward`` model.py:73-74 → nn.MultiheadAttention(batch_first=True)
    (x_q, x_kv, x_kv, key_padding_mask=pad_mask)[0] at dropout 0:
    q,k,v projections; heads are contiguous dh-chunks of E; S = (q·dh^-½) kᵀ; padded keys get
    −inf; P = softmax(S); O = P v; out = O W_oᵀ + b_o."""
    wq, wk, wv, bq, bk, bv, wo, bo = mha_weights(sd, prefix)
    b, n, e = x_q.shape
    m = x_kv.shape[1]
    dh = e // num_heads
    q = linear(x_q, wq, bq).view(b, n, num_heads, dh).transpose(1, 2)
    k = linear(x_kv, wk, bk).view(b, m, num_heads, dh).transpose(1, 2)
    v = linear(x_kv, wv, bv).view(b, m, num_heads, dh).transpose(1, 2)
    s = (q * (dh ** -0.5)) @ k.transpose(-1, -2)
    if pad_mask is not None:
        s = s.masked_fill(pad_mask[:, None, None, :], float("-inf"))
    p = torch.softmax(s, dim=-1)
    o = (p @ v).transpose(1, 2).reshape(b, n, e)
    return linear(o, wo, bo)


# --------------------------------------------------------------------------------------
# model.py building blocks
# --------------------------------------------------------------------------------------

def mlp_residual(sd: StateDict, prefix: str, x: torch.Tensor) -> torch.Tensor:
    """``Residual(mlp(C))`` model.py:20-26,47-56 at dropout 0.  Keys: ``<prefix>module.{0,1,3}``."""
    h = layer_norm(x, sd[prefix + "module.0.weight"], sd[prefix + "module.0.bias"])
    h = linear(h, sd[prefix + "module.1.weight"], sd[prefix + "module.1.bias"])
    h = gelu_erf(h)
    h = linear(h, sd[prefix + "module.3.weight"], sd[prefix + "module.3.bias"])
    return h + x


def cross_attention_residual(sd: StateDict, prefix: str, x_q: torch.Tensor, x_kv: torch.Tensor,
                             num_heads: int, pad_mask: Optional[torch.Tensor]) -> torch.Tensor:
    """``Residual(CrossAttention)`` model.py:77-99 + :54-56: q_norm / kv_norm, MHA, then add the
    *un-normalised* query (``args[0]``)."""
    q = layer_norm(x_q, sd[prefix + "module.q_norm.weight"], sd[prefix + "module.q_norm.bias"])
    kv = layer_norm(x_kv, sd[prefix + "module.kv_norm.weight"], sd[prefix + "module.kv_norm.bias"])
    return multi_head_attention(sd, prefix + "module.attention.attention.", q, kv, num_heads, pad_mask) + x_q


def self_attention_residual(sd: StateDict, prefix: str, x: torch.Tensor, num_heads: int) -> torch.Tensor:
    """``Residual(SelfAttention)`` model.py:102-116.  No mask reaches it (utils.py:10-13 passes a
    lone tensor through the Sequential)."""
    h = layer_norm(x, sd[prefix + "module.norm.weight"], sd[prefix + "module.norm.bias"])
    return multi_head_attention(sd, prefix + "module.attention.attention.", h, h, num_heads, None) + x


def cross_attention_layer(sd, prefix, x_q, x_kv, num_heads, pad_mask):
    """model.py:29-33 — Sequential(Residual(CrossAttention), Residual(mlp))."""
    x = cross_attention_residual(sd, prefix + "0.", x_q, x_kv, num_heads, pad_mask)
    return mlp_residual(sd, prefix + "1.", x)


def self_attention_block(sd, prefix, x, num_heads, num_layers):
    """model.py:36-44."""
    for s in range(num_layers):
        x = self_attention_residual(sd, f"{prefix}{s}.0.", x, num_heads)
        x = mlp_residual(sd, f"{prefix}{s}.1.", x)
    return x


def perceiver_layer(sd, prefix, x_latent, x, cross_heads, self_heads, sa_layers, pad_mask):
    """``create_perceiver_layer`` model.py:150-160."""
    x_latent = cross_attention_layer(sd, prefix + "0.", x_latent, x, cross_heads, pad_mask)
    return self_attention_block(sd, prefix + "1.", x_latent, self_heads, sa_layers)


# --------------------------------------------------------------------------------------
# adapters (adapter.py)
# --------------------------------------------------------------------------------------

def fourier_position_encoding(spatial_shape, num_bands: int) -> torch.Tensor:
    """adapter.py:53-94.  Layout for d spatial dims: [p_0..p_{d-1}, sin(π p_i f_i) for i, cos(...)
    for i]; f_i = linspace(1, size_i/2, F); association order π·(p·f) as at :86,91-92."""
    coords = [torch.linspace(-1.0, 1.0, steps=s) for s in spatial_shape]
    p = torch.stack(torch.meshgrid(*coords, indexing="ij"), dim=len(spatial_shape))
    grids = []
    for i, s in enumerate(spatial_shape):
        f = torch.linspace(1.0, s / 2.0, num_bands)
        grids.append(p[..., i:i + 1] * f[None, ...])
    enc = [p] + [torch.sin(math.pi * g) for g in grids] + [torch.cos(math.pi * g) for g in grids]
    enc = torch.cat(enc, dim=-1)
    return enc.reshape(-1, enc.shape[-1])


def image_input_adapter(sd: StateDict, prefix: str, x: torch.Tensor) -> torch.Tensor:
    """adapter.py:99-109: flatten spatial dims, broadcast the table over batch, concat."""
    b = x.shape[0]
    enc = sd[prefix + "position_encoding"]
    x = x.reshape(b, -1, x.shape[-1])
    return torch.cat([x, enc.unsqueeze(0).expand(b, -1, -1)], dim=-1)


def text_input_adapter(sd: StateDict, prefix: str, ids: torch.Tensor) -> torch.Tensor:
    """adapter.py:127-133: Embedding(ids)·√C + pos[:l]."""
    emb = sd[prefix + "text_embedding.weight"]
    l = ids.shape[1]
    return emb[ids] * math.sqrt(emb.shape[1]) + sd[prefix + "pos_encoding"][:l]


def output_adapter_linear(sd: StateDict, prefix: str, x: torch.Tensor) -> torch.Tensor:
    """adapter.py:146-149 (Classification/TextOutputAdapter): linear then squeeze(dim=1)."""
    return linear(x, sd[prefix + "linear.weight"], sd[prefix + "linear.bias"]).squeeze(dim=1)


# --------------------------------------------------------------------------------------
# encoder / decoder / whole models
# --------------------------------------------------------------------------------------

def encoder(sd: StateDict, prefix: str, x: torch.Tensor, *, kind: str, num_layers: int, cross_heads: int,
            self_heads: int, sa_layers: int, pad_mask: Optional[torch.Tensor] = None) -> torch.Tensor:
    """``PerceiverEncoder.forward`` model.py:176-189; ``layer_n`` weights are shared by the
    (num_layers-1) repeats (:186-187)."""
    if kind == "image":
        x = image_input_adapter(sd, prefix + "input_adapter.", x)
    elif kind == "text":
        x = text_input_adapter(sd, prefix + "input_adapter.", x)
    else:
        raise ValueError(kind)
    b = x.shape[0]
    lat = sd[prefix + "latent"].unsqueeze(0).expand(b, -1, -1)
    lat = perceiver_layer(sd, prefix + "layer_1.", lat, x, cross_heads, self_heads, sa_layers, pad_mask)
    for _ in range(num_layers - 1):
        lat = perceiver_layer(sd, prefix + "layer_n.", lat, x, cross_heads, self_heads, sa_layers, pad_mask)
    return lat


def decoder(sd: StateDict, prefix: str, x_latent: torch.Tensor, *, num_heads: int) -> torch.Tensor:
    """``PerceiverDecoder.forward`` model.py:229-237: learned query array, cross-attention layer
    without any mask (pad_mask is accepted and ignored), output adapter."""
    b = x_latent.shape[0]
    q = sd[prefix + "output"].unsqueeze(0).expand(b, -1, -1)
    out = cross_attention_layer(sd, prefix + "cross_attention.", q, x_latent, num_heads, None)
    return output_adapter_linear(sd, prefix + "output_adapter.", out)


def perceiver_io(sd: StateDict, x: torch.Tensor, pad_mask: Optional[torch.Tensor] = None, *, kind: str,
                 num_layers: int, cross_heads: int, self_heads: int, sa_layers: int, dec_heads: int,
                 enc_prefix: str = "0.", dec_prefix: str = "1.") -> torch.Tensor:
    """``PerceiverIO`` model.py:321-325 = Sequential(encoder, decoder) (prefixes '0.' / '1.')."""
    lat = encoder(sd, enc_prefix, x, kind=kind, num_layers=num_layers, cross_heads=cross_heads,
                  self_heads=self_heads, sa_layers=sa_layers, pad_mask=pad_mask)
    return decoder(sd, dec_prefix, lat, num_heads=dec_heads)


def text_masking(x: torch.Tensor, pad_mask: torch.Tensor, *, vocab_size: int, unk_token_id: int,
                 mask_token_id: int, num_special_tokens: int, mask_p: float = 0.15,
                 generator: Optional[torch.Generator] = None) -> Tuple[torch.Tensor, torch.Tensor]:
    """``TextMasking.forward`` model.py:265-293 (BERT 80/10/10).  Draw order — three float32
    uniforms of x's shape, then ``randint(num_special, vocab, (count,))`` — is part of the
    contract: identical device + generator state ⇒ identical (x_masked, labels).  In place on
    ``x`` like the reference (:281,286)."""
    labels = x.clone()
    is_special = (x == unk_token_id) | pad_mask
    shape, dev = x.shape, x.device
    r0 = torch.rand(shape, dtype=torch.float, device=dev, generator=generator)
    selected = (r0 < mask_p) & ~is_special
    r1 = torch.rand(shape, dtype=torch.float, device=dev, generator=generator)
    sel1 = selected & (r1 < 0.9)
    r2 = torch.rand(shape, dtype=torch.float, device=dev, generator=generator)
    sel2 = sel1 & (r2 < 1 / 9)
    x[sel1] = mask_token_id
    x[sel2] = torch.randint(num_special_tokens, vocab_size, size=(int(sel2.sum()),), device=dev,
                            generator=generator)
    labels[~selected] = -100
    return x, labels


def cross_entropy(logits: torch.Tensor, labels: torch.Tensor, ignore_index: int = -100) -> torch.Tensor:
    """nn.CrossEntropyLoss() over the class dim, mean over non-ignored targets
    (lightning.py:61,184,225-226).  ``logits`` (..., classes), ``labels`` (...)."""
    logits = logits.reshape(-1, logits.shape[-1])
    labels = labels.reshape(-1)
    valid = labels != ignore_index
    lse = torch.logsumexp(logits, dim=-1)
    picked = logits.gather(1, labels.clamp(min=0)[:, None])[:, 0]
    return ((lse - picked) * valid).sum() / valid.sum()


def mlm_forward(sd: StateDict, ids_masked: torch.Tensor, pad_mask: Optional[torch.Tensor], *, num_layers: int,
                cross_heads: int, self_heads: int, sa_layers: int, dec_heads: int) -> torch.Tensor:
    """Repaired ``PerceiverMLM.forward`` model.py:306-318 (the fork's own wiring raises,
    SURVEY.md §0): encoder → decoder → ``[:, :l, :]``; keys prefixed 'encoder.' / 'decoder.'."""
    l = ids_masked.shape[1]
    logits = perceiver_io(sd, ids_masked, pad_mask, kind="text", num_layers=num_layers, cross_heads=cross_heads,
                          self_heads=self_heads, sa_layers=sa_layers, dec_heads=dec_heads,
                          enc_prefix="encoder.", dec_prefix="decoder.")
    return logits[:, :l, :]
