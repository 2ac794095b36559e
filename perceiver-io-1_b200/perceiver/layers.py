"""Parameter holders with torch-compatible names, shapes and initialisation.

The reference builds its layers from nn.LayerNorm / nn.Linear / nn.MultiheadAttention; the
drop-in keeps their parameter names and shapes (so reference checkpoints load) and their
initialisation order (same seed -> same weights), but the forward arithmetic is the fused CUDA
path in functional.py."""
import math

import torch
import torch.nn as nn

from . import functional as F


class LayerNorm(nn.Module):
    def __init__(self, num_channels: int, eps: float = 1e-5):
        super().__init__()
        self.normalized_shape = (num_channels,)
        self.eps = eps
        self.weight = nn.Parameter(torch.ones(num_channels))
        self.bias = nn.Parameter(torch.zeros(num_channels))

    def extra_repr(self):
        return f'{self.normalized_shape[0]}, eps={self.eps}'


class Linear(nn.Module):
    def __init__(self, in_features: int, out_features: int):
        super().__init__()
        self.in_features, self.out_features = in_features, out_features
        self.weight = nn.Parameter(torch.empty(out_features, in_features))
        self.bias = nn.Parameter(torch.empty(out_features))
        # nn.Linear.reset_parameters
        nn.init.kaiming_uniform_(self.weight, a=math.sqrt(5))
        bound = 1 / math.sqrt(in_features) if in_features > 0 else 0
        nn.init.uniform_(self.bias, -bound, bound)

    def forward(self, x):
        return F.LinearFn.apply(x, self.weight, self.bias)

    def extra_repr(self):
        return f'in_features={self.in_features}, out_features={self.out_features}'


class MultiheadAttentionParams(nn.Module):
    """Parameters of torch.nn.MultiheadAttention(embed_dim, num_heads, kdim, vdim, batch_first=True):
    packed ``in_proj_weight (3E,E)`` when kdim == vdim == E, otherwise ``q/k/v_proj_weight``;
    ``in_proj_bias (3E,)``; ``out_proj.{weight,bias}``."""

    def __init__(self, embed_dim: int, num_heads: int, kdim: int, vdim: int, dropout: float = 0.0):
        super().__init__()
        if embed_dim % num_heads != 0:
            raise AssertionError('embed_dim must be divisible by num_heads')
        self.embed_dim, self.num_heads, self.kdim, self.vdim, self.dropout = embed_dim, num_heads, kdim, vdim, dropout
        self._qkv_same_embed_dim = kdim == embed_dim and vdim == embed_dim
        if self._qkv_same_embed_dim:
            self.in_proj_weight = nn.Parameter(torch.empty(3 * embed_dim, embed_dim))
        else:
            self.q_proj_weight = nn.Parameter(torch.empty(embed_dim, embed_dim))
            self.k_proj_weight = nn.Parameter(torch.empty(embed_dim, kdim))
            self.v_proj_weight = nn.Parameter(torch.empty(embed_dim, vdim))
        self.in_proj_bias = nn.Parameter(torch.empty(3 * embed_dim))
        self.out_proj = Linear(embed_dim, embed_dim)  # draws its own init first, like torch
        if self._qkv_same_embed_dim:
            nn.init.xavier_uniform_(self.in_proj_weight)
        else:
            nn.init.xavier_uniform_(self.q_proj_weight)
            nn.init.xavier_uniform_(self.k_proj_weight)
            nn.init.xavier_uniform_(self.v_proj_weight)
        with torch.no_grad():
            self.in_proj_bias.zero_()
            self.out_proj.bias.zero_()

    def packed(self):
        """(w_packed, w_q, w_k, w_v): exactly one of w_packed / the split triple is set."""
        if self._qkv_same_embed_dim:
            return self.in_proj_weight, None, None, None
        return None, self.q_proj_weight, self.k_proj_weight, self.v_proj_weight
