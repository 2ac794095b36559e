"""B200-native drop-in for the `perceiver` package of felixyu7/perceiver-io-1 (hot path only).

Put `perceiver-io-1_b200/` on PYTHONPATH ahead of the reference checkout and
`from perceiver import PerceiverIO, ...` resolves here.  All forward/backward arithmetic runs in
hand-written sm_100a CUDA kernels (libpio_b200.so, C ABI in include/pio_b200.h).
"""
from .adapter import (ClassificationOutputAdapter, ImageInputAdapter, InputAdapter, OutputAdapter,
                      SemanticSegOutputAdapter, TextInputAdapter, TextOutputAdapter)
from .lightning import LitImageClassifier, LitMaskedLanguageModel, LitTextClassifier
from .functional import cross_entropy, get_compute_mode, linear_cross_entropy, set_compute_mode
from .model import (CrossAttention, MultiHeadAttention, PerceiverDecoder, PerceiverEncoder, PerceiverIO, PerceiverMLM,
                    Residual, SelfAttention, TextMasking, cross_attention_layer, mlp, self_attention_block,
                    self_attention_layer)
from .distributed import FlatGradients, average_gradients
from .optim import FusedAdamW
from .pipeline import GraphedMLMStep, GraphedTrainStep, MaskingPrefetcher
from .utils import Sequential, freeze, predict_masked_samples
