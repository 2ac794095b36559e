"""ctypes binding of libpio_b200.so (declarations: include/pio_b200.h).

There is no CPU or torch fallback: if the library is missing the import of any compute entry
point raises, loudly.  PyTorch is used only for device memory and streams.
"""
import ctypes
import os
import threading
from ctypes import (POINTER, Structure, byref, c_char_p, c_float, c_int, c_int32, c_int64, c_longlong, c_size_t,
                    c_void_p)

import torch

_PKG_ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
LIB_PATH = os.environ.get("PIO_B200_LIB", os.path.join(_PKG_ROOT, "libpio_b200.so"))

MODE_FP32 = 0
MODE_BF16 = 1
ACC_DX = 4   # PIO_ACC_DX: the activation gradient is added into its buffer
ACC_DKV = 8  # PIO_ACC_DKV

_lib = None
_lock = threading.Lock()


class AttnWeights(Structure):
    _names = ("ln_w", "ln_b", "wq", "wk", "wv", "bq", "bk", "bv", "wo", "bo", "kv_ln_w", "kv_ln_b")
    _fields_ = [(n, c_void_p) for n in _names]


class MlpWeights(Structure):
    _names = ("ln_w", "ln_b", "w1", "b1", "w2", "b2")
    _fields_ = [(n, c_void_p) for n in _names]


class GemmDesc(Structure):
    _fields_ = [("a", c_void_p), ("b", c_void_p), ("c", c_void_p),
                ("m", c_int64), ("n", c_int64), ("k", c_int64),
                ("lda", c_int64), ("ldb", c_int64), ("ldc", c_int64),
                ("a_kmajor", c_int), ("b_kmajor", c_int),
                ("bias", c_void_p), ("residual", c_void_p), ("ldr", c_int64),
                ("aux", c_void_p), ("ldaux", c_int64),
                ("act", c_int), ("accumulate", c_int), ("rowsum_a", c_void_p),
                ("dropout_p", c_float), ("dropout_seed", ctypes.c_uint64)]


class GemmBf16Desc(Structure):
    _fields_ = [("a", c_void_p), ("b", c_void_p),
                ("m", c_int64), ("n", c_int64), ("k", c_int64), ("lda", c_int64), ("ldb", c_int64),
                ("a_kmajor", c_int), ("b_kmajor", c_int),
                ("c32", c_void_p), ("ldc32", c_int64), ("c16", c_void_p), ("ldc16", c_int64),
                ("bias", c_void_p), ("residual", c_void_p), ("ldr", c_int64),
                ("aux16", c_void_p), ("ldaux", c_int64), ("act", c_int), ("accumulate", c_int),
                ("dropout_p", c_float), ("dropout_seed", ctypes.c_uint64),
                ("dyn_rows", c_void_p), ("dyn_offset", c_int64), ("dyn_axis", c_int)]


class Dropout(Structure):
    _fields_ = [("p", c_float), ("seed", ctypes.c_uint64)]


class AttnDesc(Structure):
    _fields_ = [("q", c_void_p), ("k", c_void_p), ("v", c_void_p), ("pad", c_void_p), ("o", c_void_p), ("lse", c_void_p),
                ("q_bs", c_int64), ("q_rs", c_int64), ("k_bs", c_int64), ("k_rs", c_int64),
                ("v_bs", c_int64), ("v_rs", c_int64), ("o_bs", c_int64), ("o_rs", c_int64),
                ("batch", c_int), ("heads", c_int), ("n", c_int), ("m", c_int), ("dh", c_int),
                ("scale", c_float)]


class AttnBwdDesc(Structure):
    _fields_ = [("fwd", AttnDesc), ("d_o", c_void_p), ("dq", c_void_p), ("dk", c_void_p), ("dv", c_void_p),
                ("delta", c_void_p)]


P = c_void_p
_SIGNATURES = {
    # name: (restype, argtypes)
    "pio_last_error": (c_char_p, []),
    "pio_version": (c_int, []),
    "pio_launch_count": (c_longlong, []),
    "pio_image_encode": (c_int, [P, P, P, c_int64, c_int64, c_int, c_int, P]),
    "pio_fourier_table": (c_int, [P, POINTER(c_int), c_int, c_int, P]),
    "pio_text_embed_fwd": (c_int, [P, P, P, P, c_int64, c_int64, c_int, c_int64, c_float, P]),
    "pio_text_embed_bwd": (c_int, [P, P, P, P, c_int64, c_int64, c_int, c_int64, c_float, P]),
    "pio_text_mask_select": (c_int, [P, P, P, P, P, P, P, c_int64, c_int64, c_float, P]),
    "pio_text_mask_apply": (c_int, [P, P, P, P, P, c_int64, c_int64, c_int64, P]),
    "pio_layernorm_fwd": (c_int, [P, P, P, P, P, c_int64, c_int, c_float, P]),
    "pio_layernorm_bwd": (c_int, [P, P, P, P, P, P, P, P, c_int, c_int64, c_int, P, c_size_t, P]),
    "pio_layernorm_bwd_ws_bytes": (c_size_t, [c_int64, c_int]),
    "pio_layernorm_bwd_noalloc": (c_int, [P, P, P, P, P, P, P, P, c_int, c_int64, c_int, P]),
    "pio_gemm_f32": (c_int, [POINTER(GemmDesc), P, c_size_t, P]),
    "pio_gemm_f32_ws_bytes": (c_size_t, [c_int64, c_int64, c_int64]),
    "pio_gemm_bf16": (c_int, [POINTER(GemmBf16Desc), P, c_size_t, P]),
    "pio_gemm_bf16_ws_bytes": (c_size_t, [c_int64, c_int64, c_int64]),
    "pio_cast_bf16": (c_int, [P, P, c_int64, c_int, c_int64, c_int64, P]),
    "pio_cast_bf16_multi": (c_int, [P, c_int, P]),
    "pio_adamw_step": (c_int, [P, P, c_int, P, P]),
    "pio_adamw_tick": (c_int, [P, P]),
    "pio_flat_copy_f32": (c_int, [P, P, P, c_int, c_int, ctypes.c_float, P]),
    "pio_scale3_f32": (c_int, [P, c_int64, P, c_int64, P, c_int64, P, P]),
    "pio_latent_block_fwd_pre": (c_int, [P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, c_int, c_int, c_int, c_int,
                                     c_int, POINTER(Dropout), c_int, P, c_size_t, P, P]),
    "pio_latent_block_bwd_pre": (c_int, [P, P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, POINTER(AttnWeights),
                                     POINTER(MlpWeights), c_int, c_int, c_int, c_int, c_int, c_int, POINTER(Dropout), c_int, P,
                                     c_size_t, P, P, P]),
    "pio_cross_attend_o16_offset": (c_size_t, [c_int, c_int, c_int, c_int]),
    "pio_cross_attend_merges_dkv": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "pio_latent_block_cluster_ok": (c_int, [c_int, c_int, c_int, c_int, c_int]),
    "pio_set_side_join_deferred": (c_int, [c_int]),
    "pio_side_join": (c_int, [P]),
    "pio_logits_bf16_supported": (c_int, [c_int64, c_int64, c_int64, c_int64, c_int64, c_int64]),
    "pio_logits_bf16": (c_int, [P, c_int64, P, c_int64, P, P, c_int64, c_int64, c_int64, c_int64, P, c_int, P]),
    "pio_dw_batched": (c_int, [P, c_int, c_int64, P]),
    "pio_ln_param_grads": (c_int, [P, c_int, c_int64, c_int, P]),
    "pio_attn_fwd": (c_int, [POINTER(AttnDesc), c_int, P]),
    "pio_attn_bwd": (c_int, [POINTER(AttnBwdDesc), c_int, P]),
    "pio_attn_bwd_bf16": (c_int, [POINTER(AttnDesc), P, P, P, P, P, P]),
    "pio_cross_entropy_fwd": (c_int, [P, P, P, P, c_int64, c_int64, c_int64, c_int64, P]),
    "pio_cross_entropy_bwd": (c_int, [P, P, P, P, P, P, c_int64, c_int64, c_int64, c_int64, P]),
    "pio_mlp_saved_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "pio_mlp_ws_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "pio_mlp_fwd": (c_int, [P, P, POINTER(MlpWeights), P, c_int64, c_int, c_int, POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_mlp_bwd": (c_int, [P, P, P, POINTER(MlpWeights), P, POINTER(MlpWeights), c_int, c_int64, c_int, c_int,
                            POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_kv_project_saved_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
    "pio_kv_project_ws_bytes": (c_size_t, [c_int64, c_int, c_int, c_int]),
    "pio_kv_project_fwd": (c_int, [P, P, POINTER(AttnWeights), P, c_int64, c_int, c_int, c_int, P, c_size_t, P]),
    "pio_kv_project_bwd": (c_int, [P, P, P, POINTER(AttnWeights), P, POINTER(AttnWeights), c_int, c_int64, c_int, c_int,
                                   c_int, P, c_size_t, P]),
    "pio_cross_attend_saved_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "pio_cross_attend_ws_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "pio_cross_attend_fwd": (c_int, [P, P, P, P, POINTER(AttnWeights), P, c_int, c_int, c_int, c_int, c_int, c_int,
                                     POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_cross_attend_bwd": (c_int, [P, P, P, P, P, P, POINTER(AttnWeights), P, POINTER(AttnWeights), c_int, c_int, c_int,
                                     c_int, c_int, c_int, c_int, POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_self_attn_saved_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "pio_self_attn_ws_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int]),
    "pio_self_attn_fwd": (c_int, [P, P, POINTER(AttnWeights), P, c_int, c_int, c_int, c_int, c_int, POINTER(Dropout), c_int, P,
                                  c_size_t, P]),
    "pio_self_attn_bwd": (c_int, [P, P, P, POINTER(AttnWeights), P, POINTER(AttnWeights), c_int, c_int, c_int, c_int,
                                  c_int, c_int, POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_latent_block_saved_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "pio_latent_block_ws_bytes": (c_size_t, [c_int, c_int, c_int, c_int, c_int, c_int]),
    "pio_latent_block_fwd": (c_int, [P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, c_int, c_int, c_int, c_int,
                                     c_int, POINTER(Dropout), c_int, P, c_size_t, P]),
    "pio_latent_block_bwd": (c_int, [P, P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, POINTER(AttnWeights),
                                     POINTER(MlpWeights), c_int, c_int, c_int, c_int, c_int, c_int, POINTER(Dropout), c_int, P,
                                     c_size_t, P]),
    "pio_label_compact_ws_bytes": (c_size_t, [c_int64]),
    "pio_label_compact": (c_int, [P, c_int64, c_int64, c_int64, P, P, P, P, c_size_t, P]),
    "pio_decode_rows_supported": (c_int, [c_int, c_int, c_int]),
    "pio_decode_rows_saved_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "pio_decode_rows_ws_bytes": (c_size_t, [c_int64, c_int, c_int]),
    "pio_decode_rows_fwd": (c_int, [P, P, P, P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, c_int64, c_int, c_int, c_int, c_int,
                                    c_int, P, c_size_t, P]),
    "pio_decode_rows_bwd": (c_int, [P, P, P, P, P, P, POINTER(AttnWeights), POINTER(MlpWeights), P, POINTER(AttnWeights),
                                    POINTER(MlpWeights), c_int, c_int64, c_int, c_int, c_int, c_int, c_int, P, c_size_t, P]),
    "pio_linear_ws_bytes": (c_size_t, [c_int64, c_int, c_int64, c_int]),
    "pio_linear_fwd": (c_int, [P, P, P, P, c_int64, c_int, c_int64, c_int, P, c_size_t, P]),
    "pio_linear_ce_ws_bytes": (c_size_t, [c_int64, c_int, c_int64, c_int]),
    "pio_linear_ce_fwd_bwd": (c_int, [P, P, P, P, P, P, P, P, c_int64, c_int, c_int64, c_int64, c_int, P, c_size_t, P]),
    "pio_linear_bwd": (c_int, [P, P, P, P, P, P, c_int, c_int64, c_int, c_int64, c_int, P, c_size_t, P]),
}

EXPORTS = tuple(_SIGNATURES)


def lib():
    """The loaded library; raises RuntimeError when it has not been built."""
    global _lib
    if _lib is None:
        with _lock:
            if _lib is None:
                if not os.path.exists(LIB_PATH):
                    raise RuntimeError(
                        f"{LIB_PATH} not found: build it with `python perceiver-io-1_b200/build.py` "
                        "(there is no CPU / PyTorch fallback for the Perceiver IO kernels)")
                handle = ctypes.CDLL(LIB_PATH)
                for name, (res, args) in _SIGNATURES.items():
                    fn = getattr(handle, name)  # AttributeError if the .so is stale
                    fn.restype = res
                    fn.argtypes = args
                _lib = handle
    return _lib


def check(status: int, what: str = ""):
    if status != 0:
        msg = lib().pio_last_error().decode(errors="replace")
        raise RuntimeError(f"pio_b200 error {status}{' in ' + what if what else ''}: {msg}")


def ptr(t):
    return None if t is None else t.data_ptr()


def stream_ptr(device=None):
    """the caller's current stream on ``device``.  The library resolves per-device resources (side streams, the SM count,
    function attributes, the no-allocation scratch array) for the CURRENT device, so the tensors' device must be the
    current one: anything else is refused here, loudly, instead of pairing a stream with another device's resources."""
    if device is not None:
        idx = torch.device(device).index
        if idx is not None and idx != torch.cuda.current_device():
            raise RuntimeError(f"perceiver (B200-native): tensors live on cuda:{idx} but the current device is "
                               f"cuda:{torch.cuda.current_device()}; wrap the call in `with torch.cuda.device({idx}):`")
    return torch.cuda.current_stream(device).cuda_stream


def require_cuda(*tensors):
    for t in tensors:
        if t is not None and not t.is_cuda:
            raise RuntimeError("perceiver (B200-native): tensors must live on a CUDA device; there is no CPU path")


# ---- scratch workspace: one growing buffer per (device, stream) --------------------------------
_ws = {}


def workspace(nbytes: int, device, tag=None):
    """scratch for one operator call; `tag`: a separate buffer (for scratch that outlives the call, see
    functional.LatentBlockFn.backward)"""
    key = (device.index if device.index is not None else torch.cuda.current_device(),
           torch.cuda.current_stream(device).cuda_stream, tag)
    buf = _ws.get(key)
    if buf is None or buf.numel() < nbytes:
        buf = torch.empty(max(int(nbytes * 1.25), 1 << 20), dtype=torch.uint8, device=device)
        _ws[key] = buf
    return buf


def attn_struct(**tensors):
    s = AttnWeights()
    for k, v in tensors.items():
        setattr(s, k, v if isinstance(v, int) or v is None else v.data_ptr())
    return s


def mlp_struct(**tensors):
    s = MlpWeights()
    for k, v in tensors.items():
        setattr(s, k, v if isinstance(v, int) or v is None else v.data_ptr())
    return s
