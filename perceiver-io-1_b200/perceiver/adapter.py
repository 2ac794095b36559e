"""Input / output adapters — same constructors, properties, state_dict keys and errors as the
reference's perceiver/adapter.py; the arithmetic runs in libpio_b200.so."""
import math
from typing import Optional, Tuple

import torch
import torch.nn as nn

from . import functional as F
from .layers import Linear


class InputAdapter(nn.Module):
    def __init__(self, num_input_channels):
        super().__init__()
        self._num_input_channels = num_input_channels

    @property
    def num_input_channels(self):
        return self._num_input_channels

    def forward(self, x):
        raise NotImplementedError()


class OutputAdapter(nn.Module):
    def __init__(self, output_shape):
        super().__init__()
        self._output_shape = output_shape

    @property
    def output_shape(self):
        return self._output_shape

    def forward(self, x):
        raise NotImplementedError()


def fourier_position_table(spatial_shape, num_frequency_bands):
    """Host-side (init-time) Fourier table with the reference's exact operation order
    (adapter.py:53-94) so the persisted buffer is identical: positions linspace(-1,1) per axis
    ('ij' mesh), f_i = linspace(1, size_i/2, F), then [p, sin(pi*(p_i f_i))..., cos(...)...]."""
    axes = [torch.linspace(-1.0, 1.0, steps=s) for s in spatial_shape]
    pos = torch.stack(torch.meshgrid(*axes, indexing='ij'), dim=len(spatial_shape))
    scaled = []
    for i, size in enumerate(spatial_shape):
        freqs = torch.linspace(1.0, size / 2.0, num_frequency_bands)
        scaled.append(pos[..., i:i + 1] * freqs[None, ...])
    parts = [pos] + [torch.sin(math.pi * s) for s in scaled] + [torch.cos(math.pi * s) for s in scaled]
    table = torch.cat(parts, dim=-1)
    return table.reshape(-1, table.shape[-1])


class ImageInputAdapter(InputAdapter):
    def __init__(self, image_shape: Tuple[int, ...], num_frequency_bands: int):
        *self.spatial_shape, num_image_channels = image_shape
        self.image_shape = image_shape
        self.num_frequency_bands = num_frequency_bands
        super().__init__(num_input_channels=num_image_channels + self._num_position_encoding_channels())
        self.register_buffer('position_encoding', fourier_position_table(self.spatial_shape, num_frequency_bands))

    def _num_position_encoding_channels(self, include_positions: bool = True) -> int:
        return len(self.spatial_shape) * (2 * self.num_frequency_bands + include_positions)

    def regenerate_on_device(self):
        """Rebuild the table with the on-device generator (pio_fourier_table); useful for very large
        inputs (e.g. 512x512 -> 262144 x 130).  Differs from the host table by sin/cos ulps."""
        from . import _native as N
        dev = self.position_encoding.device
        N.require_cuda(self.position_encoding)
        import ctypes
        shape = (ctypes.c_int * len(self.spatial_shape))(*self.spatial_shape)
        out = torch.empty_like(self.position_encoding)
        N.check(N.lib().pio_fourier_table(out.data_ptr(), shape, len(self.spatial_shape), self.num_frequency_bands,
                                          N.stream_ptr(dev)), "pio_fourier_table")
        self.position_encoding.copy_(out)
        return self

    def forward(self, x):
        b, *d = x.shape
        if tuple(d) != tuple(self.image_shape):
            raise ValueError(f'Input image shape {tuple(d)} different from required shape {self.image_shape}')
        return F.ImageEncodeFn.apply(x.reshape(b, self.position_encoding.shape[0], d[-1]), self.position_encoding)


class TextInputAdapter(InputAdapter):
    def __init__(self, vocab_size: int, max_seq_len: int, num_input_channels: int):
        super().__init__(num_input_channels=num_input_channels)
        self.text_embedding = nn.Embedding(vocab_size, num_input_channels)  # parameter holder only
        self.pos_encoding = nn.Parameter(torch.empty(max_seq_len, num_input_channels))
        self.scale = math.sqrt(num_input_channels)
        with torch.no_grad():
            self.text_embedding.weight.data.uniform_(-0.1, 0.1)
            self.pos_encoding.uniform_(-0.5, 0.5)

    def forward(self, x):
        b, l = x.shape
        if l > self.pos_encoding.shape[0]:
            raise ValueError(f'sequence length {l} exceeds max_seq_len {self.pos_encoding.shape[0]}')
        return F.TextEmbedFn.apply(x, self.text_embedding.weight, self.pos_encoding, self.scale)


class ClassificationOutputAdapter(OutputAdapter):
    def __init__(self, num_classes: int, num_outputs: int = 1, num_output_channels: Optional[int] = None):
        if num_output_channels is None:
            num_output_channels = num_classes
        super().__init__(output_shape=(num_outputs, num_output_channels))
        self.linear = Linear(num_output_channels, num_classes)

    def forward(self, x):
        return self.linear(x).squeeze(dim=1)


class SemanticSegOutputAdapter(OutputAdapter):
    """Identity forward with an unused linear, as in the reference (adapter.py:151-164)."""

    def __init__(self, num_classes: int, num_outputs: int = 1, num_output_channels: Optional[int] = None):
        if num_output_channels is None:
            num_output_channels = num_classes
        super().__init__(output_shape=(num_outputs, num_output_channels))
        self.linear = Linear(num_output_channels, num_classes)

    def forward(self, x):
        return x


class TextOutputAdapter(ClassificationOutputAdapter):
    def __init__(self, vocab_size: int, max_seq_len: int, num_output_channels: Optional[int] = None):
        super().__init__(num_classes=vocab_size, num_outputs=max_seq_len, num_output_channels=num_output_channels)
