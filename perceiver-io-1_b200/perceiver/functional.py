"""torch.autograd glue over the C ABI: one Function per fused operator of libpio_b200.so.

Everything numerical happens in the CUDA kernels; torch provides storage, streams and the
autograd graph.  Saved-for-backward state is an opaque byte buffer whose layout only the
library knows (pio_*_saved_bytes).
"""
import ctypes
import os
import threading
import weakref
from ctypes import byref

import torch
from torch.autograd.function import once_differentiable

from . import _native as N

_MODE = {"mode": N.MODE_FP32}


def set_compute_mode(mode):
    """'fp32' (FFMA kernels, rel 1e-4 parity mode) or 'bf16' (tcgen05 tensor cores, rel 1e-2)."""
    m = {"fp32": N.MODE_FP32, "bf16": N.MODE_BF16, N.MODE_FP32: N.MODE_FP32, N.MODE_BF16: N.MODE_BF16}.get(mode)
    if m is None:
        raise ValueError(f"unknown compute mode {mode!r}")
    _MODE["mode"] = m


def get_compute_mode():
    return _MODE["mode"]


_drop_calls = [0]


def next_dropout_seed():
    """A fresh 64-bit seed for one operator call, derived from torch's seed and a call counter (so
    torch.manual_seed makes dropout reproducible).  Operators use seed .. seed + 4*layers for their sites."""
    _drop_calls[0] += 1
    return (torch.initial_seed() * 0x9E3779B97F4A7C15 + _drop_calls[0] * 0x632BE59BD9B4E019) & 0xFFFFFFFFFFFFFFFF


def _drop_struct(p, seed):
    """ctypes pio_dropout (or None when off)"""
    if not p:
        return None
    return N.Dropout(float(p), int(seed))


def _dref(d):
    return byref(d) if d is not None else None


def _c(t, dtype=torch.float32):
    """contiguous CUDA tensor of the expected dtype (no silent device moves)."""
    N.require_cuda(t)
    if t.dtype != dtype:
        raise TypeError(f"expected {dtype}, got {t.dtype}")
    return t if t.is_contiguous() else t.contiguous()


def _saved(nbytes, device):
    return torch.empty(int(nbytes), dtype=torch.uint8, device=device)


def invalidate_weight_cache(module=None):
    """Drop bf16 weight shadows (only needed after writing a frozen parameter through ``param.data``
    outside training; trainable parameters are re-cast on every training forward anyway)."""
    if module is not None:
        _TABLES.pop(module, None)
        for p in module.parameters():
            if hasattr(p, "_pio_w16"):
                del p._pio_w16


def refresh_weight_shadows(module):
    """Re-cast, IN PLACE, every bf16 weight shadow the last training forward of ``module`` used (one launch).  For
    code that writes trainable parameters behind the optimizer's back (``param.copy_`` / ``load_state_dict``) while
    a captured CUDA graph holds the shadow addresses: the graph does not re-cast them itself when the optimizer is
    `perceiver.FusedAdamW`, which writes the shadows together with the parameters."""
    tab = _TABLES.get(module)
    if tab is None:
        return False
    _, table, weights = tab
    N.check(N.lib().pio_cast_bf16_multi(table.data_ptr(), len(weights), N.stream_ptr(table.device)), "pio_cast_bf16_multi")
    for w in weights:
        hit = getattr(w, "_pio_w16", None)
        if hit is not None:
            w._pio_w16 = (w._version, w.data_ptr(), hit[2], hit[3], hit[4])
    return True


# bf16 weight shadows.  A shadow is valid for ONE training forward: optimizers such as
# torch.optim.AdamW(fused=True) update parameters without moving their version counters, so "version
# unchanged" proves nothing for a trainable parameter.  Model-level forwards open a `shadow_scope`; the
# first one records which matrices were shadowed, every later one re-casts all of them with a single
# pio_cast_bf16_multi launch (in place: the buffers keep their addresses, so captured graphs stay valid).
# Frozen parameters and no-grad forwards keep the version-keyed cache.
_SH = {"epoch": 0, "depth": 0, "seen": None, "grad": True}
_TABLES = weakref.WeakKeyDictionary()  # top-level module -> (key, device table, weights)


def _trainable(w):
    """may an optimizer have touched `w` since its shadow was made?  Grad mode is read where the scope was
    opened (inside autograd.Function.forward it is always off); outside any scope assume yes."""
    return w.requires_grad and (_SH["grad"] if _SH["depth"] > 0 else True)


def _cast_one(w, out):
    rows, cols = w.shape
    ld = (cols + 7) // 8 * 8
    if out is None or out.shape != (rows, ld) or out.device != w.device:
        out = torch.empty(rows, ld, dtype=torch.bfloat16, device=w.device)
    N.check(N.lib().pio_cast_bf16(w.data_ptr(), out.data_ptr(), rows, cols, cols, ld, N.stream_ptr(w.device)), "pio_cast_bf16")
    return out


def _w16(w):
    """bf16 shadow of a 2-D fp32 weight, row pitch padded to 8 elements (TMA); lives on the parameter
    object.  Entry = (version, data_ptr, shadow, epoch, made_in_training)."""
    hit = getattr(w, "_pio_w16", None)
    train = _trainable(w)
    epoch = _SH["epoch"] if _SH["depth"] > 0 else None
    same = hit is not None and hit[0] == w._version and hit[1] == w.data_ptr()
    if same and ((train and epoch is not None and hit[3] == epoch) or (not train and not hit[4])):
        out = hit[2]
    else:
        out = _cast_one(w, hit[2] if hit is not None else None)
        try:
            w._pio_w16 = (w._version, w.data_ptr(), out, epoch, train)
        except Exception:
            pass
    if train and _SH["seen"] is not None:
        _SH["seen"][id(w)] = w
    return out


def _refresh_shadows(module, epoch):
    """one launch re-casting every matrix the previous training forward of `module` shadowed"""
    tab = _TABLES.get(module)
    if tab is None:
        return
    _, table, weights = tab
    for w in weights:
        hit = getattr(w, "_pio_w16", None)
        if hit is None or hit[1] != w.data_ptr() or not w.requires_grad or w.device != table.device:
            _TABLES.pop(module, None)  # parameters were replaced / moved / frozen: start over
            return
    if all(getattr(w, "_pio_fresh16", False) and w._pio_w16[0] == w._version for w in weights):
        pass  # perceiver.optim.FusedAdamW wrote every shadow together with the parameter: nothing to cast
    else:
        N.check(N.lib().pio_cast_bf16_multi(table.data_ptr(), len(weights), N.stream_ptr(table.device)), "pio_cast_bf16_multi")
    for w in weights:
        w._pio_fresh16 = False
        w._pio_w16 = (w._version, w.data_ptr(), w._pio_w16[2], epoch, True)


def _remember_shadows(module, seen):
    weights = [w for w in seen.values() if getattr(w, "_pio_w16", None) is not None]
    if not weights:
        return
    key = tuple((id(w), w.data_ptr(), w._pio_w16[2].data_ptr()) for w in weights)
    tab = _TABLES.get(module)
    if tab is not None and tab[0] == key:
        return
    dev = weights[0].device
    if any(w.device != dev for w in weights) or torch.cuda.is_current_stream_capturing():
        return
    rows = []
    for w in weights:
        out = w._pio_w16[2]
        rows.append([w.data_ptr(), out.data_ptr(), w.shape[0], w.shape[1] | (out.shape[1] << 32)])  # pio_cast_item
    table = torch.tensor(rows, dtype=torch.int64).to(dev)
    _TABLES[module] = (key, table, weights)


class shadow_scope:
    """Opened by the model-level forwards (PerceiverIO, PerceiverMLM, encoder, decoder); nests."""

    def __init__(self, module):
        self.module = module

    def __enter__(self):
        _SH["depth"] += 1
        if _SH["depth"] == 1:
            _SH["grad"] = torch.is_grad_enabled()
            _forget_pending()  # (a backward pass that died half-way never ran its clean-up callback)
        if _SH["depth"] == 1 and _MODE["mode"] == N.MODE_BF16 and _SH["grad"]:
            _SH["epoch"] += 1
            _SH["seen"] = {}
            _refresh_shadows(self.module, _SH["epoch"])
        return self

    def __exit__(self, *exc):
        _SH["depth"] -= 1
        if _SH["depth"] == 0:
            seen, _SH["seen"] = _SH["seen"], None
            if seen and exc[0] is None:
                _remember_shadows(self.module, seen)
        return False


def _mat(w, mode):
    """the matrix operand the kernels read in this mode: the fp32 parameter or its bf16 shadow"""
    if w is None or mode == N.MODE_FP32:
        return w
    return _w16(w)


# Parameters applied more than once in a forward (the shared `layer_n` of PerceiverEncoder, reference
# model.py:186-187; the packed in-projection that KVProjectFn and CrossAttendFn both read) get one gradient buffer
# per backward pass: the first backward that reaches such a parameter allocates it and hands it to autograd, later
# ones add into the same memory (the kernels accumulate with atomics anyway) and report None — so autograd has
# nothing left to sum, which was one small add kernel per shared tensor per step.
#   * Only parameters whose gradient autograd actually asked for (ctx.needs_input_grad) are registered: a frozen
#     parameter's buffer is dropped by autograd as soon as the backward returns, and a later operator must not be
#     handed its address.  (The kernels still need somewhere to write, so a frozen parameter gets scratch space in
#     the slab; its gradient is computed and discarded.)
#   * The slab every buffer of a call is carved from is kept alive until the end of the backward pass
#     (_PASS_SLABS, cleared by the engine's end-of-pass callback), so a registered address can never point at
#     memory the caching allocator has handed to someone else.  Holding the slab does not stop autograd from
#     adopting the per-parameter views as .grad without a copy (that test looks at the view's own reference count).
#   * Known limit, by construction of "report None": a parameter shared between one of these operators and a
#     foreign torch op that ALSO produces a gradient for it in the same pass relies on autograd's own summation;
#     the Perceiver models here never do that.
# The tables are shared by all threads under a lock (autograd runs a device's backward on a worker thread, and the
# end-of-pass callback may run on another one); keys carry the device, and a pass whose entries were cleared early by
# another device's callback just allocates fresh buffers again, which autograd then sums the ordinary way.
class _PassState:
    def __init__(self):
        self.lock = threading.RLock()
        self.pending = {}  # (data_ptr, numel, device) of a parameter -> device address of its gradient buffer
        self.slabs = []    # keeps the memory behind those addresses alive for the pass
        self.acts = {}     # activation with two consumers -> the gradient buffer its first backward returned (_act_pending)
        self._tls = threading.local()

    @property
    def reused(self):  # between _grads_like and _out of ONE backward call: same thread
        r = getattr(self._tls, "reused", None)
        if r is None:
            r = self._tls.reused = set()
        return r


_PASS = _PassState()

# Flat gradient destinations (perceiver.distributed.FlatGradients): a parameter registered here has a permanent,
# pre-zeroed home for its gradient inside one flat buffer (p.grad is a view of it).  The backward operators add
# straight into that memory and report None, so there is no per-operator slab to zero, nothing for autograd to
# accumulate, nothing to flatten before the data-parallel all-reduce — and the owner is told after every operator
# which parameters were written, which is how it knows the moment a bucket is complete and its all-reduce can start
# while the rest of the backward pass is still running.
_DST = {}          # (data_ptr, numel, device) of a parameter -> (device address of its gradient, owner)
_TOUCH_LOG = None  # FlatGradients' profiling pass: parameter keys in the order the backward pass wrote them


def _pkey(p):
    return (p.data_ptr(), p.numel(), p.device)


def _touched_list():
    t = getattr(_PASS._tls, "touched", None)
    if t is None:
        t = _PASS._tls.touched = []
    return t


def _direct_target(p):
    """Backward of an operator that writes a parameter gradient itself: the address of the parameter's registered
    destination (the caller then writes / adds there and returns None for it), or None.  Counts as this pass's
    write of the parameter either way (profiling), so call it exactly once per gradient produced."""
    k = _pkey(p)
    if _TOUCH_LOG is not None:
        _TOUCH_LOG.append(k)
    hit = _DST.get(k)
    if hit is None:
        return None
    _touched_list().append((hit[1], k))
    return hit[0]


def _dst_address(p):
    """the registered destination of p's gradient, or None — without counting a write (see LinearCrossEntropyFn)"""
    hit = _DST.get(_pkey(p))
    return None if hit is None else hit[0]


def _flush_deferred():
    """An operator left parameter-gradient kernels running on the side streams (LatentBlockFn.backward): make the current
    stream wait for them, release what they read, and only now tell the owners that those destinations are written."""
    d = getattr(_PASS._tls, "deferred", None)
    if not d:
        return
    _PASS._tls.deferred = None
    for dev, entries, _keepalive in d:
        N.check(N.lib().pio_side_join(N.stream_ptr(dev)), "pio_side_join")
        for owner, k in entries:
            owner._written(k)


def _defer(dev, keepalive):
    """Move this operator's write reports (and the buffers its side-stream kernels still read) to the deferred list.
    False if nobody would flush them (not inside a backward pass): the caller must not defer then."""
    if getattr(_PASS._tls, "deferred", None) is None:
        _PASS._tls.deferred = []
    t = _touched_list()
    _PASS._tls.deferred.append((dev, list(t), keepalive))
    t.clear()
    _PASS._tls.just_deferred = True  # (the _after_op that ends THIS operator must not flush it again)
    return True


def _defer_probe():
    """True inside a backward pass (where an end-of-pass callback can be queued): registers the flush once per pass."""
    if getattr(_PASS._tls, "deferred", None) is not None:
        return True
    try:
        torch.autograd.Variable._execution_engine.queue_callback(_flush_deferred)
    except RuntimeError:
        return False
    _PASS._tls.deferred = []
    return True


def _after_op():
    """end of a backward operator: tell the owners which destinations are now written"""
    if getattr(_PASS._tls, "just_deferred", False):
        _PASS._tls.just_deferred = False
        return
    _flush_deferred()
    t = _touched_list()
    if t:
        done = list(t)
        t.clear()
        for owner, k in done:
            owner._written(k)


class _Alias:
    """stands in for a gradient tensor that autograd already holds: just its address"""

    def __init__(self, address):
        self._address = address

    def data_ptr(self):
        return self._address

    def element_size(self):
        return 4


def _forget_pending():
    with _PASS.lock:
        _PASS.pending.clear()
        _PASS.slabs.clear()
        _PASS.acts.clear()


# An ACTIVATION that two of these operators consume (the embedded input that both key/value projections read, the
# projected kv that the two applications of the shared cross-attention layer read, the latent array that goes both into
# the cross-attention and — as the residual — into the block behind it) gets the same treatment as a shared parameter:
# the first backward that produces its gradient returns the buffer to autograd and publishes it here, the second adds its
# contribution into that memory inside the kernel that computes it (PIO_ACC_DX / PIO_ACC_DKV) and reports None, so autograd
# has nothing to sum — which was one full read-read-write pass over the activation per pair.
#   * "Same activation" means the same autograd output (grad_fn and output number, both held by the entry), the same
#     memory and the same shape; leaves and anything `_c` had to copy never match and take autograd's own summation.
#   * The entry keeps the activation and its gradient alive until the end of the pass, so neither address can be
#     handed to another tensor while a consumer may still look it up.
#   * autograd stores the first gradient of an input as it is (no copy), which is what makes adding into it later
#     the same as summing; bf16 mode only (the fp32 operators refuse the flags).
#   * That holds while autograd's input buffer for the activation IS the published tensor.  A further consumer outside
#     this protocol (a skip connection a user adds) makes autograd sum into another tensor as soon as its gradient
#     arrives, and a share added to the published buffer after that would be lost.  So a later producer merges only
#     while the engine still holds the published tensor — its reference count says so (this table's Python object + the
#     engine's buffer = 2; autograd never adds in place into a tensor someone else holds, and a sum formed AFTER a merge
#     contains it) — and otherwise returns its gradient the ordinary way.  Without `Tensor._use_count` nothing is merged.
_ACT_STATS = {"merged": 0}  # how many activation gradients were added in place (tests)


def _act_key(x):
    return (x.data_ptr(), tuple(x.shape), x.dtype, x.device)


def _act_pending(x):
    """the gradient buffer another operator of this backward pass already returned for activation ``x``, or None"""
    if x.grad_fn is None or os.environ.get("PIO_NO_GRAD_MERGE") is not None:
        return None
    with _PASS.lock:
        e = _PASS.acts.get(_act_key(x))
    if e is not None and e[0] is x.grad_fn and e[1] == x.output_nr:
        try:
            held = e[2]._use_count() >= 2  # (1: autograd has replaced its buffer by a sum with a foreign consumer's gradient)
        except AttributeError:
            held = False
        if held:
            _ACT_STATS["merged"] += 1
            return e[2]
    return None


def _act_publish(x, dx):
    """first producer of ``x``'s gradient in this pass: later producers may add into ``dx`` (no-op outside a backward
    pass, where nothing would clear the entry)"""
    if x.grad_fn is None or dx is None or os.environ.get("PIO_NO_GRAD_MERGE") is not None:
        return
    with _PASS.lock:
        if not _PASS.acts:
            try:
                torch.autograd.Variable._execution_engine.queue_callback(_forget_pending)
            except RuntimeError:
                return
        _PASS.acts[_act_key(x)] = (x.grad_fn, x.output_nr, dx, x)


def _grads_like(params, mode, needs=None):
    """Gradient buffers shaped like ``params`` (None stays None) and the `accumulate` flag to pass.
    fp32 mode: uninitialised buffers, overwritten (0).  bf16 mode: views into one zero-filled slab so the
    kernels can add into them with atomics (2) and skip their reduction passes; a parameter that already got a
    buffer earlier in this backward pass gets that one again (return the result through `_out`).
    ``needs[i]`` (ctx.needs_input_grad of the parameter) False: scratch that is neither registered nor returned."""
    if needs is None:
        needs = [True] * len(params)
    if mode == N.MODE_FP32:
        return [torch.empty_like(p) if p is not None else None for p in params], 0
    st = _PASS
    st.reused.clear()
    with st.lock:
        return _grads_like_locked(st, params, needs)


def _grads_like_locked(st, params, needs):
    keys = [None if p is None else _pkey(p) for p in params]
    kinds = []  # per parameter: None | "dst" | "pending" | "fresh"
    for p, k, need in zip(params, keys, needs):
        if p is None:
            kinds.append(None)
        elif need and k in _DST:
            kinds.append("dst")
        elif k in st.pending:
            kinds.append("pending")
        else:
            kinds.append("fresh")
    fresh = [p for p, kind in zip(params, kinds) if kind == "fresh"]
    offs, total = [], 0
    for p in fresh:
        offs.append(total)
        total += (p.numel() + 3) // 4 * 4
    slab = torch.zeros(total, dtype=torch.float32, device=fresh[0].device) if fresh else None
    share = bool(st.slabs)  # a non-empty list means this pass already registered its end-of-pass callback
    if fresh and not share:
        try:
            torch.autograd.Variable._execution_engine.queue_callback(_forget_pending)
            share = True
        except RuntimeError:  # not inside a backward pass (direct call): nothing to share across calls
            share = False
    if share and slab is not None:
        st.slabs.append(slab)
    it = iter(offs)
    out = []
    touched = _touched_list()
    for p, k, need, kind in zip(params, keys, needs, kinds):
        if kind is None:
            out.append(None)
            continue
        if need and _TOUCH_LOG is not None:
            _TOUCH_LOG.append(k)
        if kind == "dst":
            addr, owner = _DST[k]
            g = _Alias(addr)
            st.reused.add(id(g))
            touched.append((owner, k))
        elif kind == "pending":
            g = _Alias(st.pending[k])
            st.reused.add(id(g))
        else:
            o = next(it)
            g = slab[o:o + p.numel()].view(p.shape)
            if not need:
                st.reused.add(id(g))  # scratch: reported as None
            elif share:
                st.pending[k] = g.data_ptr()
        out.append(g)
    return out, 2


def _out(*grads):
    """what a backward returns: buffers autograd already holds, and scratch of frozen parameters (see _grads_like),
    are reported as None"""
    reused = _PASS.reused
    if not reused:
        return grads
    res = tuple(None if (g is not None and id(g) in reused) else g for g in grads)
    reused.clear()
    _after_op()
    return res


def _mha_ptrs(w_packed, w_q, w_k, w_v, bias, e):
    """Device addresses of (wq, wk, wv, bq, bk, bv) for packed (3E,E) or split storage; matrix
    weights may be fp32 or bf16 shadows (the packed row pitch is E in both)."""
    if w_packed is not None:
        base = w_packed.data_ptr()
        esz = w_packed.element_size()
        wq, wk, wv = base, base + esz * e * e, base + 2 * esz * e * e
    else:
        wq = w_q.data_ptr() if w_q is not None else None
        wk = w_k.data_ptr() if w_k is not None else None
        wv = w_v.data_ptr() if w_v is not None else None
    b = bias.data_ptr()
    return wq, wk, wv, b, b + 4 * e, b + 8 * e


# ------------------------------------------------------------------------------------------
# adapters
# ------------------------------------------------------------------------------------------

class ImageEncodeFn(torch.autograd.Function):
    """ImageInputAdapter.forward (reference adapter.py:99-109)."""

    @staticmethod
    def forward(ctx, x, pos_enc):
        x = _c(x)
        pos_enc = _c(pos_enc)
        b = x.shape[0]
        c_img = x.shape[-1]
        m, c_pos = pos_enc.shape
        out = torch.empty(b, m, c_img + c_pos, dtype=torch.float32, device=x.device)
        N.check(N.lib().pio_image_encode(x.data_ptr(), pos_enc.data_ptr(), out.data_ptr(), b, m, c_img, c_pos,
                                         N.stream_ptr(x.device)), "pio_image_encode")
        ctx.x_shape = x.shape
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        dx = dout[..., :ctx.x_shape[-1]].reshape(ctx.x_shape) if ctx.needs_input_grad[0] else None
        return dx, None


class TextEmbedFn(torch.autograd.Function):
    """TextInputAdapter.forward (reference adapter.py:127-133)."""

    @staticmethod
    def forward(ctx, ids, emb, pos, scale):
        ids = _c(ids, torch.int64)
        emb = _c(emb)
        pos = _c(pos)
        b, l = ids.shape
        vocab, c = emb.shape
        out = torch.empty(b, l, c, dtype=torch.float32, device=emb.device)
        N.check(N.lib().pio_text_embed_fwd(ids.data_ptr(), emb.data_ptr(), pos.data_ptr(), out.data_ptr(), b, l, c, vocab,
                                           float(scale), N.stream_ptr(emb.device)), "pio_text_embed_fwd")
        ctx.save_for_backward(ids, emb, pos)
        ctx.meta = (vocab, c, pos.shape[0], float(scale))
        return out

    @staticmethod
    @once_differentiable
    def backward(ctx, dout):
        ids, emb, pos = ctx.saved_tensors
        vocab, c, max_len, scale = ctx.meta
        dout = _c(dout)
        b, l = ids.shape
        # a registered flat destination (FlatGradients) is zero at this point and takes the scatter-add directly
        a_emb = _direct_target(emb) if ctx.needs_input_grad[1] else None
        a_pos = _direct_target(pos) if ctx.needs_input_grad[2] else None
        demb = dpos = None
        if ctx.needs_input_grad[1] and a_emb is None:
            demb = torch.zeros(vocab, c, dtype=torch.float32, device=dout.device)
            a_emb = demb.data_ptr()
        if ctx.needs_input_grad[2] and a_pos is None:
            dpos = torch.zeros(max_len, c, dtype=torch.float32, device=dout.device)
            a_pos = dpos.data_ptr()
        N.check(N.lib().pio_text_embed_bwd(ids.data_ptr(), dout.data_ptr(), a_emb, a_pos, b, l, c, vocab, scale,
                                           N.stream_ptr(dout.device)), "pio_text_embed_bwd")
        _after_op()
        return None, demb, dpos, None


class BroadcastRowsFn(torch.autograd.Function):
    """(K, C) parameter (first ``rows`` rows) -> (B, rows, C): the learned latent array / decoder query array repeated
    over the batch (reference model.py:183, :235 `repeat(self.latent, '... -> b ...')`).  Backward: the sum over the
    batch, written straight into the parameter's flat gradient destination when one is registered (FlatGradients)."""

    @staticmethod
    def forward(ctx, param, batch, rows):
        ctx.save_for_backward(param)
        ctx.rows = rows
        return param[:rows].unsqueeze(0).expand(batch, -1, -1).contiguous()

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        (param,) = ctx.saved_tensors
        rows = ctx.rows
        hit = _DST.get(_pkey(param))
        if hit is not None and param.grad is not None and param.grad.data_ptr() == hit[0]:
            _direct_target(param)  # counts the write
            torch.sum(dy, dim=0, out=param.grad[:rows])  # (rows beyond `rows` stay zero: the buffer was cleared)
            _after_op()
            return None, None, None
        if _TOUCH_LOG is not None:
            _TOUCH_LOG.append(_pkey(param))
        g = dy.sum(dim=0)
        if rows != param.shape[0]:
            full = torch.zeros_like(param)
            full[:rows] = g
            g = full
        return g, None, None


def text_mask(ids, pad_mask, vocab_size, unk_token_id, mask_token_id, num_special_tokens, mask_p, generator=None):
    """TextMasking.forward (reference model.py:265-293), in place on ``ids``; returns labels.

    The four RNG draws stay torch calls in the reference's order (three float32 uniforms, then a
    randint whose length is the number of random-replacement positions), so the same device and
    generator state give bit-identical results; the select / scatter logic is two kernels."""
    N.require_cuda(ids, pad_mask)
    if ids.dtype != torch.int64 or not ids.is_contiguous():
        raise TypeError("ids must be a contiguous int64 tensor (masked in place)")
    dev = ids.device
    n = ids.numel()
    pad = None
    if pad_mask is not None:
        if pad_mask.dtype != torch.bool:
            raise TypeError("pad_mask must be a bool tensor")
        pad = pad_mask.expand_as(ids).contiguous()
    r0 = torch.rand(ids.shape, dtype=torch.float, device=dev, generator=generator)
    r1 = torch.rand(ids.shape, dtype=torch.float, device=dev, generator=generator)
    r2 = torch.rand(ids.shape, dtype=torch.float, device=dev, generator=generator)
    nblk = max(1, (n + 1023) // 1024)
    flags = torch.empty(max(n, 1), dtype=torch.uint8, device=dev)
    counts = torch.zeros(nblk, dtype=torch.int32, device=dev)
    st = N.stream_ptr(dev)
    N.check(N.lib().pio_text_mask_select(ids.data_ptr(), N.ptr(pad), r0.data_ptr(), r1.data_ptr(), r2.data_ptr(),
                                         flags.data_ptr(), counts.data_ptr(), n, int(unk_token_id), float(mask_p), st),
            "pio_text_mask_select")
    n_random = int(counts.sum().item())  # the reference's own host sync: is_selected_2.sum() (model.py:288)
    rnd = torch.randint(num_special_tokens, vocab_size, size=(n_random,), device=dev, generator=generator)
    labels = torch.empty_like(ids)
    N.check(N.lib().pio_text_mask_apply(ids.data_ptr(), labels.data_ptr(), flags.data_ptr(), counts.data_ptr(),
                                        rnd.data_ptr() if n_random else None, n_random, n, int(mask_token_id), st),
            "pio_text_mask_apply")
    return labels


# ------------------------------------------------------------------------------------------
# fused operators
# ------------------------------------------------------------------------------------------

def _mlp_operands(params, mode):
    ln_w, ln_b, w1, b1, w2, b2 = params
    return (ln_w, ln_b, _mat(w1, mode), b1, _mat(w2, mode), b2)


class MlpFn(torch.autograd.Function):
    """Residual(mlp(C)) (reference model.py:20-26,47-56)."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, w1, b1, w2, b2, residual, drop_p=0.0, drop_seed=0):
        x = _c(x)
        params = [_c(p) for p in (ln_w, ln_b, w1, b1, w2, b2)]
        c = x.shape[-1]
        rows = x.numel() // c
        mode = get_compute_mode()
        L = N.lib()
        saved = _saved(L.pio_mlp_saved_bytes(rows, c, mode), x.device)
        ws = N.workspace(L.pio_mlp_ws_bytes(rows, c, mode), x.device)
        y = torch.empty_like(x)
        ops = _mlp_operands(params, mode)
        w = N.mlp_struct(**dict(zip(N.MlpWeights._names, ops)))
        drop = _drop_struct(drop_p, drop_seed)
        N.check(L.pio_mlp_fwd(x.data_ptr(), y.data_ptr(), byref(w), saved.data_ptr(), rows, c, int(residual), _dref(drop), mode,
                              ws.data_ptr(), ws.numel(), N.stream_ptr(x.device)), "pio_mlp_fwd")
        ctx.save_for_backward(x, *params)
        ctx.wstruct, ctx.keep = w, ops
        ctx.saved_buf = saved
        ctx.meta = (rows, c, int(residual), mode, drop)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, *params = ctx.saved_tensors
        rows, c, residual, mode, drop = ctx.meta
        dy = _c(dy)
        L = N.lib()
        grads, accf = _grads_like(params, mode, ctx.needs_input_grad[1:7])
        dx = torch.empty_like(x)
        ws = N.workspace(L.pio_mlp_ws_bytes(rows, c, mode), x.device)
        w = ctx.wstruct
        g = N.mlp_struct(**dict(zip(N.MlpWeights._names, grads)))
        N.check(L.pio_mlp_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), byref(w), ctx.saved_buf.data_ptr(), byref(g), accf,
                              rows, c, residual, _dref(drop), mode, ws.data_ptr(), ws.numel(), N.stream_ptr(x.device)),
                "pio_mlp_bwd")
        ctx.saved_buf = None
        return _out(dx, *grads, None, None, None)


class KVProjectFn(torch.autograd.Function):
    """kv_norm + key/value in-projection of CrossAttention (reference model.py:90,98 and the k/v
    half of nn.MultiheadAttention's in-projection, :74).  Output kv (B, M, 2E) = [K | V]."""

    @staticmethod
    def forward(ctx, x_kv, ln_w, ln_b, bias, w_packed, w_k, w_v, e):
        x_kv = _c(x_kv)
        ln_w, ln_b, bias = _c(ln_w), _c(ln_b), _c(bias)
        w_packed = _c(w_packed) if w_packed is not None else None
        w_k = _c(w_k) if w_k is not None else None
        w_v = _c(w_v) if w_v is not None else None
        b, m, ckv = x_kv.shape
        rows = b * m
        mode = get_compute_mode()
        L = N.lib()
        mats = (_mat(w_packed, mode), _mat(w_k, mode), _mat(w_v, mode))
        _, wk, wv, _, bk, bv = _mha_ptrs(mats[0], None, mats[1], mats[2], bias, e)
        w = N.attn_struct(wk=wk, wv=wv, bk=bk, bv=bv, kv_ln_w=ln_w, kv_ln_b=ln_b)
        ctx.wstruct, ctx.keep = w, mats
        saved = _saved(L.pio_kv_project_saved_bytes(rows, ckv, e, mode), x_kv.device)
        ws = N.workspace(L.pio_kv_project_ws_bytes(rows, ckv, e, mode), x_kv.device)
        kv = torch.empty(b, m, 2 * e, dtype=torch.float32 if mode == N.MODE_FP32 else torch.bfloat16, device=x_kv.device)
        N.check(L.pio_kv_project_fwd(x_kv.data_ptr(), kv.data_ptr(), byref(w), saved.data_ptr(), rows, ckv, e, mode,
                                     ws.data_ptr(), ws.numel(), N.stream_ptr(x_kv.device)), "pio_kv_project_fwd")
        ctx.save_for_backward(x_kv, ln_w, ln_b, bias, w_packed, w_k, w_v)
        ctx.saved_buf = saved
        ctx.meta = (rows, ckv, e, mode)
        return kv

    @staticmethod
    @once_differentiable
    def backward(ctx, dkv):
        x_kv, ln_w, ln_b, bias, w_packed, w_k, w_v = ctx.saved_tensors
        rows, ckv, e, mode = ctx.meta
        dkv = _c(dkv, torch.float32 if mode == N.MODE_FP32 else torch.bfloat16)
        L = N.lib()
        dev = x_kv.device
        if mode == N.MODE_FP32:
            dln_w, dln_b = torch.empty_like(ln_w), torch.empty_like(ln_b)
            dbias = torch.zeros_like(bias)
            dwp = torch.zeros_like(w_packed) if w_packed is not None else None
            dwk = torch.empty_like(w_k) if w_k is not None else None
            dwv = torch.empty_like(w_v) if w_v is not None else None
            accf = 0
        else:
            (dln_w, dln_b, dbias, dwp, dwk, dwv), accf = _grads_like((ln_w, ln_b, bias, w_packed, w_k, w_v), mode,
                                                                    ctx.needs_input_grad[1:7])
        dx = merged = None
        if ctx.needs_input_grad[0]:
            merged = _act_pending(x_kv) if mode == N.MODE_BF16 else None  # the other key/value projection of the same input
            dx = torch.empty_like(x_kv) if merged is None else merged
        _, gwk, gwv, _, gbk, gbv = _mha_ptrs(dwp, None, dwk, dwv, dbias, e)
        w = ctx.wstruct
        g = N.attn_struct(wk=gwk, wv=gwv, bk=gbk, bv=gbv, kv_ln_w=dln_w, kv_ln_b=dln_b)
        ws = N.workspace(L.pio_kv_project_ws_bytes(rows, ckv, e, mode), dev)
        N.check(L.pio_kv_project_bwd(dkv.data_ptr(), x_kv.data_ptr(), N.ptr(dx), byref(w), ctx.saved_buf.data_ptr(),
                                     byref(g), accf | (N.ACC_DX if merged is not None else 0), rows, ckv, e, mode,
                                     ws.data_ptr(), ws.numel(), N.stream_ptr(dev)),
                "pio_kv_project_bwd")
        ctx.saved_buf = None
        if merged is not None:
            dx = None
        elif mode == N.MODE_BF16:
            _act_publish(x_kv, dx)
        return _out(dx, dln_w, dln_b, dbias, dwp, dwk, dwv, None)


class CrossAttendFn(torch.autograd.Function):
    """Query side of Residual(CrossAttention) (reference model.py:77-99, :54-56): q_norm,
    q in-projection, attention over the projected kv, out-projection, + x_q.
    ``residual=2`` (bf16 mode): stop after the attention and return its output (B, N, E) as bf16 — the out-projection,
    its residual and their gradients then belong to the consumer (LatentBlockFn's first layer); ``wo`` / ``bo`` are None."""

    @staticmethod
    def forward(ctx, x_q, kv, pad, ln_w, ln_b, bias, w_packed, w_q, wo, bo, heads, residual, drop_p=0.0, drop_seed=0):
        mode = get_compute_mode()
        x_q, kv = _c(x_q), _c(kv, torch.float32 if mode == N.MODE_FP32 else torch.bfloat16)
        ln_w, ln_b, bias = _c(ln_w), _c(ln_b), _c(bias)
        deferred = int(residual) == 2
        if deferred:
            assert mode == N.MODE_BF16 and wo is None and bo is None
        else:
            wo, bo = _c(wo), _c(bo)
        w_packed = _c(w_packed) if w_packed is not None else None
        w_q = _c(w_q) if w_q is not None else None
        if pad is not None:
            pad = _c(pad, torch.bool)
        b, n, e = x_q.shape
        m = kv.shape[1]
        mode = get_compute_mode()
        L = N.lib()
        mats = (_mat(w_packed, mode), _mat(w_q, mode), _mat(wo, mode))
        wq, _, _, bq, _, _ = _mha_ptrs(mats[0], mats[1], None, None, bias, e)
        w = N.attn_struct(ln_w=ln_w, ln_b=ln_b, wq=wq, bq=bq, wo=mats[2], bo=bo)
        ctx.wstruct, ctx.keep = w, mats
        saved = _saved(L.pio_cross_attend_saved_bytes(b, n, m, e, heads, mode), x_q.device)
        ws = N.workspace(L.pio_cross_attend_ws_bytes(b, n, m, e, heads, mode), x_q.device)
        y = None if deferred else torch.empty_like(x_q)
        drop = _drop_struct(drop_p, drop_seed)
        N.check(L.pio_cross_attend_fwd(x_q.data_ptr(), kv.data_ptr(), N.ptr(pad), N.ptr(y), byref(w), saved.data_ptr(),
                                       b, n, m, e, heads, int(residual), _dref(drop), mode, ws.data_ptr(), ws.numel(),
                                       N.stream_ptr(x_q.device)), "pio_cross_attend_fwd")
        ctx.save_for_backward(x_q, kv, pad, ln_w, ln_b, bias, w_packed, w_q, wo, bo)
        ctx.saved_buf = saved
        ctx.meta = (b, n, m, e, heads, int(residual), mode, drop)
        if deferred:  # the attention output lives in the saved buffer
            off = int(L.pio_cross_attend_o16_offset(b, n, e, heads))
            return saved[off:off + b * n * e * 2].view(torch.bfloat16).view(b, n, e)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x_q, kv, pad, ln_w, ln_b, bias, w_packed, w_q, wo, bo = ctx.saved_tensors
        b, n, m, e, heads, residual, mode, drop = ctx.meta
        dy = _c(dy, torch.bfloat16 if residual == 2 else torch.float32)  # (deferred out-projection: the gradient of the attention output)
        L = N.lib()
        dev = x_q.device
        if mode == N.MODE_FP32:
            dln_w, dln_b = torch.empty_like(ln_w), torch.empty_like(ln_b)
            dbias = torch.zeros_like(bias)
            dwp = torch.zeros_like(w_packed) if w_packed is not None else None
            dwq = torch.empty_like(w_q) if w_q is not None else None
            dwo, dbo = torch.empty_like(wo), torch.empty_like(bo)
            accf = 0
        else:
            (dln_w, dln_b, dbias, dwp, dwq, dwo, dbo), accf = _grads_like((ln_w, ln_b, bias, w_packed, w_q, wo, bo), mode,
                                                                         ctx.needs_input_grad[3:10])
        # x_q also feeds the block behind a deferred out-projection (its residual), kv the other application of a shared layer
        dx_m = _act_pending(x_q) if residual == 2 else None
        dkv_m = _act_pending(kv) if mode == N.MODE_BF16 and bool(L.pio_cross_attend_merges_dkv(b, n, m, e, heads)) else None
        dx = torch.empty_like(x_q) if dx_m is None else dx_m
        dkv = torch.empty_like(kv) if dkv_m is None else dkv_m
        accf |= (N.ACC_DX if dx_m is not None else 0) | (N.ACC_DKV if dkv_m is not None else 0)
        gwq, _, _, gbq, _, _ = _mha_ptrs(dwp, dwq, None, None, dbias, e)
        w = ctx.wstruct
        g = N.attn_struct(ln_w=dln_w, ln_b=dln_b, wq=gwq, bq=gbq, wo=dwo, bo=dbo)
        ws = N.workspace(L.pio_cross_attend_ws_bytes(b, n, m, e, heads, mode), dev)
        N.check(L.pio_cross_attend_bwd(dy.data_ptr(), x_q.data_ptr(), kv.data_ptr(), N.ptr(pad), dx.data_ptr(),
                                       dkv.data_ptr(), byref(w), ctx.saved_buf.data_ptr(), byref(g), accf, b, n, m, e, heads,
                                       residual, _dref(drop), mode, ws.data_ptr(), ws.numel(), N.stream_ptr(dev)),
                "pio_cross_attend_bwd")
        ctx.saved_buf = None
        if dx_m is not None:
            dx = None
        elif residual == 2:
            _act_publish(x_q, dx)
        if dkv_m is not None:
            dkv = None
        elif mode == N.MODE_BF16:
            _act_publish(kv, dkv)
        return _out(dx, dkv, None, dln_w, dln_b, dbias, dwp, dwq, dwo, dbo, None, None, None, None)


_SA_NAMES = ("ln_w", "ln_b", "in_w", "in_b", "wo", "bo")
_LAYER_PARAMS = 12  # 6 attention + 6 mlp tensors per latent layer


def _sa_struct(ln_w, ln_b, in_w, in_b, wo, bo, e, mode=N.MODE_FP32):
    """``mode`` selects fp32 parameters or their bf16 shadows for the two matrices (gradient structs
    always pass MODE_FP32)."""
    in_m, wo_m = _mat(in_w, mode), _mat(wo, mode)
    wq, wk, wv, bq, bk, bv = _mha_ptrs(in_m, None, None, None, in_b, e)
    return dict(ln_w=ln_w, ln_b=ln_b, wq=wq, wk=wk, wv=wv, bq=bq, bk=bk, bv=bv, wo=wo_m, bo=bo, _keep=(in_m, wo_m))


class SelfAttnFn(torch.autograd.Function):
    """Residual(SelfAttention) (reference model.py:102-116, :54-56), one layer."""

    @staticmethod
    def forward(ctx, x, ln_w, ln_b, in_w, in_b, wo, bo, heads, residual, drop_p=0.0, drop_seed=0):
        x = _c(x)
        params = [_c(p) for p in (ln_w, ln_b, in_w, in_b, wo, bo)]
        b, n, c = x.shape
        mode = get_compute_mode()
        L = N.lib()
        fields = _sa_struct(*params, c, mode)
        ctx.keep = fields.pop("_keep")
        w = N.attn_struct(**fields)
        ctx.wstruct = w
        saved = _saved(L.pio_self_attn_saved_bytes(b, n, c, heads, mode), x.device)
        ws = N.workspace(L.pio_self_attn_ws_bytes(b, n, c, heads, mode), x.device)
        y = torch.empty_like(x)
        drop = _drop_struct(drop_p, drop_seed)
        N.check(L.pio_self_attn_fwd(x.data_ptr(), y.data_ptr(), byref(w), saved.data_ptr(), b, n, c, heads, int(residual),
                                    _dref(drop), mode, ws.data_ptr(), ws.numel(), N.stream_ptr(x.device)), "pio_self_attn_fwd")
        ctx.save_for_backward(x, *params)
        ctx.saved_buf = saved
        ctx.meta = (b, n, c, heads, int(residual), mode, drop)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, *params = ctx.saved_tensors
        b, n, c, heads, residual, mode, drop = ctx.meta
        dy = _c(dy)
        L = N.lib()
        grads, accf = _grads_like(params, mode, ctx.needs_input_grad[1:7])
        dx = torch.empty_like(x)
        w = ctx.wstruct
        gfields = _sa_struct(*grads, c)
        gfields.pop("_keep")
        g = N.attn_struct(**gfields)
        ws = N.workspace(L.pio_self_attn_ws_bytes(b, n, c, heads, mode), x.device)
        N.check(L.pio_self_attn_bwd(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), byref(w), ctx.saved_buf.data_ptr(), byref(g),
                                    accf, b, n, c, heads, residual, _dref(drop), mode, ws.data_ptr(), ws.numel(),
                                    N.stream_ptr(x.device)),
                "pio_self_attn_bwd")
        ctx.saved_buf = None
        return _out(dx, *grads, None, None, None, None)


class LatentBlockFn(torch.autograd.Function):
    """self_attention_block (reference model.py:36-44): S x [Residual(SelfAttention),
    Residual(mlp)] in one call.  ``params`` = 12 tensors per layer: ln_w, ln_b, in_proj_weight,
    in_proj_bias, out_proj.weight, out_proj.bias, then the mlp's ln_w, ln_b, w1, b1, w2, b2."""

    @staticmethod
    def _structs(params, layers, c, mode=N.MODE_FP32):
        attn = (N.AttnWeights * layers)()
        mlp = (N.MlpWeights * layers)()
        keep = []
        for s in range(layers):
            p = params[s * _LAYER_PARAMS:(s + 1) * _LAYER_PARAMS]
            if p[2] is not None:  # (a first layer without attention of its own has no in-projection: wq stays null)
                fields = _sa_struct(*p[:6], c, mode)
                keep.append(fields.pop("_keep"))
                for k, v in fields.items():
                    setattr(attn[s], k, v if isinstance(v, int) or v is None else v.data_ptr())
            elif p[4] is not None:  # ... but it may carry the out-projection of the cross-attention layer in front
                wo_m = _mat(p[4], mode)
                keep.append((wo_m,))
                attn[s].wo = wo_m.data_ptr()
                attn[s].bo = p[5].data_ptr()
            ops = _mlp_operands(p[6:], mode)
            keep.append(ops)
            for k, v in zip(N.MlpWeights._names, ops):
                setattr(mlp[s], k, v.data_ptr())
        return attn, mlp, keep

    @staticmethod
    def forward(ctx, x, o16, heads, drop_p, drop_seed, *params):
        x = _c(x)
        o16 = None if o16 is None else _c(o16, torch.bfloat16)  # attention output whose out-projection the first layer applies
        params = [None if p is None else _c(p) for p in params]
        layers = len(params) // _LAYER_PARAMS
        assert layers * _LAYER_PARAMS == len(params) and layers >= 1
        b, n, c = x.shape
        mode = get_compute_mode()
        L = N.lib()
        attn, mlp, keep = LatentBlockFn._structs(params, layers, c, mode)
        ctx.wstructs, ctx.keep = (attn, mlp), keep
        saved = _saved(L.pio_latent_block_saved_bytes(layers, b, n, c, heads, mode), x.device)
        ws = N.workspace(L.pio_latent_block_ws_bytes(layers, b, n, c, heads, mode), x.device)
        y = torch.empty_like(x)
        drop = _drop_struct(drop_p, drop_seed)
        N.check(L.pio_latent_block_fwd_pre(x.data_ptr(), y.data_ptr(), attn, mlp, saved.data_ptr(), layers, b, n, c, heads,
                                           _dref(drop), mode, ws.data_ptr(), ws.numel(), N.stream_ptr(x.device), N.ptr(o16)),
                "pio_latent_block_fwd")
        ctx.save_for_backward(x, o16, *params)
        ctx.saved_buf = saved
        ctx.meta = (layers, b, n, c, heads, mode, drop)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, o16, *params = ctx.saved_tensors
        layers, b, n, c, heads, mode, drop = ctx.meta
        dy = _c(dy)
        L = N.lib()
        _flush_deferred()  # (an earlier block's deferred kernels read the workspace this call is about to reuse)
        needs = ctx.needs_input_grad[5:]
        d_o16 = None if o16 is None else torch.empty_like(o16)
        grads, accf = _grads_like(params, mode, needs)
        dx = torch.empty_like(x)
        attn, mlp = ctx.wstructs
        attn_g, mlp_g, _ = LatentBlockFn._structs(grads, layers, c)
        # When every parameter gradient goes straight to a registered destination (FlatGradients), nothing reads them
        # until the owner is told: the operator may then leave its parameter-gradient kernels running on the side streams
        # while the next backward operator starts (pio_set_side_join_deferred).  Its scratch, the saved state and the
        # inputs stay alive until the join.
        defer = (mode == N.MODE_BF16 and accf == 2 and os.environ.get("PIO_NO_DEFERRED_JOIN") is None and
                 all(_dst_address(p) is not None for p, need in zip(params, needs) if need and p is not None))
        ws = N.workspace(L.pio_latent_block_ws_bytes(layers, b, n, c, heads, mode), x.device, "latent_bwd" if defer else None)
        if defer:
            defer = _defer_probe()
        if defer:
            L.pio_set_side_join_deferred(1)
        try:
            N.check(L.pio_latent_block_bwd_pre(dy.data_ptr(), x.data_ptr(), dx.data_ptr(), attn, mlp, ctx.saved_buf.data_ptr(), attn_g,
                                               mlp_g, accf, layers, b, n, c, heads, _dref(drop), mode, ws.data_ptr(), ws.numel(),
                                               N.stream_ptr(x.device), N.ptr(o16), N.ptr(d_o16)), "pio_latent_block_bwd")
        finally:
            if defer:
                L.pio_set_side_join_deferred(0)
        if defer:
            _defer(x.device, (ws, ctx.saved_buf, x, o16, dy, params, grads))
        ctx.saved_buf = None
        if o16 is not None and mode == N.MODE_BF16:
            _act_publish(x, dx)  # the cross-attention in front reads x too and adds its share into dx
        return _out(dx, d_o16, None, None, None, *grads)


def label_compact(labels, classes, ignore_index=-100):
    """Device-side compaction of the labelled positions of a masked-LM batch: -> (idx int32 (R,), labels_c int64 (R,),
    count int32 (1,)) with R = labels.numel(); idx[i] = flat position of the i-th labelled pair (stable order),
    labels_c[i] its label for i < count and ``ignore_index`` beyond.  The count stays on the device."""
    labels = _c(labels, torch.int64)
    rows = labels.numel()
    dev = labels.device
    idx = torch.empty(rows, dtype=torch.int32, device=dev)
    labels_c = torch.empty(rows, dtype=torch.int64, device=dev)
    count = torch.empty(1, dtype=torch.int32, device=dev)
    L = N.lib()
    ws = N.workspace(L.pio_label_compact_ws_bytes(rows), dev)
    N.check(L.pio_label_compact(labels.data_ptr(), rows, int(classes), int(ignore_index), idx.data_ptr(), labels_c.data_ptr(),
                                count.data_ptr(), ws.data_ptr(), ws.numel(), N.stream_ptr(dev)), "pio_label_compact")
    return idx, labels_c, count


def decode_rows_supported(n, e, heads):
    return get_compute_mode() == N.MODE_BF16 and bool(N.lib().pio_decode_rows_supported(int(n), int(e), int(heads)))


class DecodeRowsFn(torch.autograd.Function):
    """PerceiverDecoder.cross_attention (reference model.py:203-206, :229-237) for the labelled (sample, position) pairs
    only: -> (B*l, E) whose first ``count`` rows are the decoder outputs of the pairs ``idx`` (the rest is never read).
    bf16 mode, dropout 0.  Inputs after ``l``: q_norm weight / bias, in_proj_bias, packed in_proj_weight or q_proj_weight,
    out_proj weight / bias, heads, then the six tensors of the layer's mlp."""

    @staticmethod
    def forward(ctx, queries, kv, idx, count, l, ln_w, ln_b, bias, w_packed, w_q, wo, bo, heads, m_ln_w, m_ln_b, w1, b1, w2, b2):
        queries, kv = _c(queries), _c(kv, torch.bfloat16)
        ln_w, ln_b, bias, wo, bo = _c(ln_w), _c(ln_b), _c(bias), _c(wo), _c(bo)
        w_packed = _c(w_packed) if w_packed is not None else None
        w_q = _c(w_q) if w_q is not None else None
        mlp_params = [_c(p) for p in (m_ln_w, m_ln_b, w1, b1, w2, b2)]
        b, n, e2 = kv.shape
        e = e2 // 2
        max_rows = b * l
        if idx.numel() != max_rows or queries.shape[1] != e or queries.shape[0] < l:
            raise ValueError("DecodeRowsFn: idx / queries do not match the batch")
        mode = N.MODE_BF16
        L = N.lib()
        dev = kv.device
        mats = (_mat(w_packed, mode), _mat(w_q, mode), _mat(wo, mode))
        wq, _, _, bq, _, _ = _mha_ptrs(mats[0], mats[1], None, None, bias, e)
        aw = N.attn_struct(ln_w=ln_w, ln_b=ln_b, wq=wq, bq=bq, wo=mats[2], bo=bo)
        mops = _mlp_operands(mlp_params, mode)
        mw = N.mlp_struct(**dict(zip(N.MlpWeights._names, mops)))
        saved = _saved(L.pio_decode_rows_saved_bytes(max_rows, e, heads), dev)
        ws = N.workspace(L.pio_decode_rows_ws_bytes(max_rows, e, heads), dev)
        y = torch.empty(max_rows, e, dtype=torch.float32, device=dev)
        N.check(L.pio_decode_rows_fwd(queries.data_ptr(), kv.data_ptr(), idx.data_ptr(), count.data_ptr(), y.data_ptr(), byref(aw),
                                      byref(mw), saved.data_ptr(), max_rows, l, b, n, e, heads, ws.data_ptr(), ws.numel(),
                                      N.stream_ptr(dev)), "pio_decode_rows_fwd")
        ctx.save_for_backward(queries, kv, idx, count, ln_w, ln_b, bias, w_packed, w_q, wo, bo, *mlp_params)
        ctx.structs, ctx.keep = (aw, mw), (mats, mops)
        ctx.saved_buf = saved
        ctx.meta = (max_rows, l, b, n, e, heads)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        queries, kv, idx, count, ln_w, ln_b, bias, w_packed, w_q, wo, bo, *mlp_params = ctx.saved_tensors
        max_rows, l, b, n, e, heads = ctx.meta
        dy = _c(dy)
        L = N.lib()
        dev = kv.device
        need = ctx.needs_input_grad
        params = (queries, ln_w, ln_b, bias, w_packed, w_q, wo, bo, *mlp_params)
        needs = (need[0],) + tuple(need[5:12]) + tuple(need[13:19])
        grads, accf = _grads_like(params, N.MODE_BF16, needs)
        dqueries, dln_w, dln_b, dbias, dwp, dwq, dwo, dbo, *mg = grads
        dkv = torch.empty_like(kv)
        gwq, _, _, gbq, _, _ = _mha_ptrs(dwp, dwq, None, None, dbias, e)
        ag = N.attn_struct(ln_w=dln_w, ln_b=dln_b, wq=gwq, bq=gbq, wo=dwo, bo=dbo)
        mgs = N.mlp_struct(**dict(zip(N.MlpWeights._names, mg)))
        aw, mw = ctx.structs
        ws = N.workspace(L.pio_decode_rows_ws_bytes(max_rows, e, heads), dev)
        N.check(L.pio_decode_rows_bwd(dy.data_ptr(), kv.data_ptr(), idx.data_ptr(), count.data_ptr(), dqueries.data_ptr(),
                                      dkv.data_ptr(), byref(aw), byref(mw), ctx.saved_buf.data_ptr(), byref(ag), byref(mgs), accf,
                                      max_rows, l, b, n, e, heads, ws.data_ptr(), ws.numel(), N.stream_ptr(dev)),
                "pio_decode_rows_bwd")
        ctx.saved_buf = None
        return _out(dqueries, dkv, None, None, None, dln_w, dln_b, dbias, dwp, dwq, dwo, dbo, None, *mg)


class LinearFn(torch.autograd.Function):
    """nn.Linear of the output adapters (reference adapter.py:146-149)."""

    @staticmethod
    def forward(ctx, x, w, b):
        x, w, b = _c(x), _c(w), _c(b)
        c_out, c_in = w.shape
        rows = x.numel() // c_in
        mode = get_compute_mode()
        L = N.lib()
        ws = N.workspace(L.pio_linear_ws_bytes(rows, c_in, c_out, mode), x.device)
        y = torch.empty(*x.shape[:-1], c_out, dtype=torch.float32, device=x.device)
        wm = _mat(w, mode)
        ctx.wm = wm
        N.check(L.pio_linear_fwd(x.data_ptr(), wm.data_ptr(), b.data_ptr(), y.data_ptr(), rows, c_in, c_out, mode,
                                 ws.data_ptr(), ws.numel(), N.stream_ptr(x.device)), "pio_linear_fwd")
        ctx.save_for_backward(x, w)
        ctx.bias = b
        ctx.meta = (rows, c_in, c_out, mode)
        return y

    @staticmethod
    @once_differentiable
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        rows, c_in, c_out, mode = ctx.meta
        dy = _c(dy)
        L = N.lib()
        dx = torch.empty_like(x) if ctx.needs_input_grad[0] else None
        a_w = _direct_target(w)  # (the operator overwrites: a parameter with a flat destination is used once per pass)
        a_b = _direct_target(ctx.bias)
        dw = db = None
        if a_w is None:
            dw = torch.empty_like(w)
            a_w = dw.data_ptr()
        if a_b is None:
            db = torch.empty(c_out, dtype=torch.float32, device=x.device)
            a_b = db.data_ptr()
        ws = N.workspace(L.pio_linear_ws_bytes(rows, c_in, c_out, mode), x.device)
        N.check(L.pio_linear_bwd(dy.data_ptr(), x.data_ptr(), ctx.wm.data_ptr(), N.ptr(dx), a_w, a_b, 0, rows,
                                 c_in, c_out, mode, ws.data_ptr(), ws.numel(), N.stream_ptr(x.device)), "pio_linear_bwd")
        _after_op()
        return dx, dw, db


class CrossEntropyFn(torch.autograd.Function):
    """nn.CrossEntropyLoss(ignore_index=-100), mean reduction (reference lightning.py:61,184,
    225-226).  logits (..., classes) with the class dim LAST (no 'b m c -> b c m' transpose)."""

    @staticmethod
    def forward(ctx, logits, labels, ignore_index):
        logits = _c(logits)
        labels = _c(labels, torch.int64)
        classes = logits.shape[-1]
        rows = logits.numel() // classes
        if labels.numel() != rows:
            raise ValueError(f"labels shape {tuple(labels.shape)} does not match logits {tuple(logits.shape)}")
        dev = logits.device
        out3 = torch.empty(3, dtype=torch.float32, device=dev)
        row_lse = torch.empty(2 * max(rows, 1), dtype=torch.float32, device=dev)
        N.check(N.lib().pio_cross_entropy_fwd(logits.data_ptr(), labels.data_ptr(), out3.data_ptr(), row_lse.data_ptr(), rows,
                                              classes, classes, int(ignore_index), N.stream_ptr(dev)),
                "pio_cross_entropy_fwd")
        ctx.save_for_backward(logits, labels, out3, row_lse)
        ctx.meta = (rows, classes, int(ignore_index))
        return out3[0].clone()

    @staticmethod
    @once_differentiable
    def backward(ctx, dloss):
        logits, labels, out3, row_lse = ctx.saved_tensors
        rows, classes, ignore = ctx.meta
        dloss = dloss.to(torch.float32).reshape(1).contiguous()
        dlogits = torch.empty_like(logits)
        N.check(N.lib().pio_cross_entropy_bwd(logits.data_ptr(), labels.data_ptr(), out3.data_ptr(), row_lse.data_ptr(),
                                              dloss.data_ptr(), dlogits.data_ptr(), rows, classes, classes, ignore,
                                              N.stream_ptr(logits.device)), "pio_cross_entropy_bwd")
        return dlogits, None, None


class LinearCrossEntropyFn(torch.autograd.Function):
    """Output-adapter linear + cross-entropy in one call (bf16 mode): the logits are produced and
    consumed chunk by chunk inside L2, and the gradients of x, W and b are formed in the same pass
    (for d(loss) = 1; backward applies the incoming d(loss) with one launch that is a no-op when it is 1, which is
    what loss.backward() passes).  SURVEY.md section 8f row 1.  With flat gradient destinations registered
    (perceiver.FlatGradients) dW and db are written straight into them by the forward."""

    @staticmethod
    def forward(ctx, x, w, b, labels, ignore_index):
        x, w, b = _c(x), _c(w), _c(b)
        labels = _c(labels, torch.int64)
        classes, c_in = w.shape
        rows = x.numel() // c_in
        if labels.numel() != rows:
            raise ValueError(f"labels shape {tuple(labels.shape)} does not match inputs {tuple(x.shape)}")
        mode = get_compute_mode()
        L = N.lib()
        dev = x.device
        out3 = torch.empty(3, dtype=torch.float32, device=dev)
        dx = torch.empty_like(x)
        a_w, a_b = _dst_address(w), _dst_address(b)
        dw = torch.empty_like(w) if a_w is None else None
        db = torch.empty_like(b) if a_b is None else None
        ws = N.workspace(L.pio_linear_ce_ws_bytes(rows, c_in, classes, mode), dev)
        wm = _mat(w, mode)
        N.check(L.pio_linear_ce_fwd_bwd(x.data_ptr(), wm.data_ptr(), b.data_ptr(), labels.data_ptr(), out3.data_ptr(),
                                        dx.data_ptr(), a_w if dw is None else dw.data_ptr(), a_b if db is None else db.data_ptr(),
                                        rows, c_in, classes, int(ignore_index), mode,
                                        ws.data_ptr(), ws.numel(), N.stream_ptr(dev)), "pio_linear_ce_fwd_bwd")
        ctx.save_for_backward(dx, dw, db, w, b)
        ctx.sizes = (w.numel(), b.numel())
        return out3[0].clone()

    @staticmethod
    @once_differentiable
    def backward(ctx, dloss):
        dx, dw, db, w, b = ctx.saved_tensors
        nw, nb = ctx.sizes
        dloss = _c(dloss.reshape(1).to(torch.float32))
        t_w, t_b = _direct_target(w), _direct_target(b)
        if (t_w is None) != (dw is not None) or (t_b is None) != (db is not None):
            raise RuntimeError("LinearCrossEntropyFn: flat gradient destinations changed between forward and backward")
        a_w = t_w if dw is None else dw.data_ptr()
        a_b = t_b if db is None else db.data_ptr()
        N.check(N.lib().pio_scale3_f32(dx.data_ptr(), dx.numel(), a_w, nw, a_b, nb, dloss.data_ptr(), N.stream_ptr(dx.device)),
                "pio_scale3_f32")
        _after_op()
        return dx, dw, db, None, None


def linear_cross_entropy(x, weight, bias, labels, ignore_index=-100):
    """loss = CE(x @ weight.T + bias, labels), mean over labels != ignore_index.  Fused in bf16 mode;
    composed from the linear and cross-entropy operators in fp32 mode."""
    if get_compute_mode() == N.MODE_BF16 and torch.is_grad_enabled():
        return LinearCrossEntropyFn.apply(x, weight, bias, labels, ignore_index)
    return CrossEntropyFn.apply(LinearFn.apply(x, weight, bias), labels, ignore_index)


def cross_entropy(logits, labels, ignore_index=-100):
    return CrossEntropyFn.apply(logits, labels, ignore_index)
