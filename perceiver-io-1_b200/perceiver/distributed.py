"""Data-parallel plumbing (the only parallelism of the reference: Lightning DDP,
`strategy: ddp_find_unused_parameters_false`, scripts/trainer.yaml:47).

The hot path shards by batch only, so the single exchange per step is the mean of the parameter
gradients over the ranks.  One process per GPU; torch.distributed (NCCL over NVLink on GPUs, gloo in
the CPU tests) carries the all-reduce.
"""
from typing import Iterable, List, Optional

import torch
import torch.distributed as dist


def flatten_gradients(params: List[torch.nn.Parameter]) -> torch.Tensor:
    """One contiguous fp32 vector of all gradients (missing gradients count as zeros)."""
    chunks = [(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in params]
    return torch.cat(chunks) if chunks else torch.empty(0)


_FLAT_PLANS = {}  # id-tuple of the parameters -> flat buffer, chunk table, tensor table (+ the pointers it was built for)
_CHUNK = 16384    # PIO_ADAMW_CHUNK


def _native_plan(params):
    """launch plan of pio_flat_copy_f32 for these gradients, or None when they are not all contiguous fp32 CUDA tensors"""
    import numpy as np
    from . import _native as N
    grads = [p.grad for p in params]
    dev = grads[0].device
    if any(g is None or not g.is_cuda or g.dtype != torch.float32 or not g.is_contiguous() or g.device != dev for g in grads):
        return None
    ident = tuple((id(p), p.data_ptr(), p.numel()) for p in params)  # (ids alone can be reused after a model is freed)
    plan = _FLAT_PLANS.get(ident)
    if plan is None:
        offs, total, chunks = [], 0, []
        for ti, g in enumerate(grads):
            offs.append(total)
            n = g.numel()
            for off in range(0, n, _CHUNK):
                chunks.append((ti | (min(_CHUNK, n - off) << 32), off))
            total += (n + 3) // 4 * 4
        plan = dict(flat=torch.empty(total, dtype=torch.float32, device=dev), offs=offs, n_chunks=len(chunks),
                    chunks=torch.from_numpy(np.asarray(chunks, dtype=np.int64).reshape(-1, 2)).to(dev),
                    table=torch.empty(len(grads), 2, dtype=torch.int64, device=dev), ptrs=None)
        _FLAT_PLANS[ident] = plan
    ptrs = tuple(g.data_ptr() for g in grads)
    if ptrs != plan["ptrs"]:  # eager training hands out new gradient buffers every step; a replayed graph does not
        # synchronous copy from pageable memory (a reused pinned staging buffer could be rewritten while in flight)
        plan["table"].copy_(torch.from_numpy(np.asarray(list(zip(ptrs, plan["offs"])), dtype=np.int64)))
        plan["ptrs"] = ptrs
    plan["lib"], plan["stream"] = N.lib(), N.stream_ptr(dev)
    return plan


def average_gradients(params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None) -> None:
    """In-place mean of every parameter gradient across the ranks of ``group`` with ONE flat all-reduce
    (the gradient set of the MLM config is 45 MB: bandwidth is irrelevant on NVLink, launch count is not).
    Equivalent to what DistributedDataParallel leaves in ``.grad``; a no-op without an initialised
    process group or with a single rank.  On the GPU the flatten / unflatten around the collective are one
    launch each (pio_flat_copy_f32; torch.cat over ~300 tensors took 470 us per step)."""
    params = [p for p in params if p.requires_grad]
    if not params or not dist.is_available() or not dist.is_initialized():
        return
    world = dist.get_world_size(group)
    if world == 1:
        return
    nccl = dist.get_backend(group) == "nccl"
    plan = _native_plan(params) if all(p.grad is not None for p in params) and params[0].is_cuda else None
    if plan is not None:
        from . import _native as N
        L, st, flat = plan["lib"], plan["stream"], plan["flat"]
        N.check(L.pio_flat_copy_f32(flat.data_ptr(), plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], 1, 1.0, st),
                "pio_flat_copy_f32")
        dist.all_reduce(flat, op=dist.ReduceOp.AVG if nccl else dist.ReduceOp.SUM, group=group)
        N.check(L.pio_flat_copy_f32(flat.data_ptr(), plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], 0,
                                    1.0 if nccl else 1.0 / world, st), "pio_flat_copy_f32")
        return
    flat = flatten_gradients(params)
    if flat.is_cuda and nccl:
        dist.all_reduce(flat, op=dist.ReduceOp.AVG, group=group)  # NCCL averages in the collective: one pass less
    else:
        dist.all_reduce(flat, op=dist.ReduceOp.SUM, group=group)
        flat.div_(world)
    offset = 0
    views = []
    for p in params:
        n = p.numel()
        views.append(flat[offset:offset + n].view_as(p))
        offset += n
    for p, v in zip(params, views):
        if p.grad is None:
            p.grad = v.clone()
    torch._foreach_copy_([p.grad for p in params], views)


class FlatGradients:
    """Gradients of ``params`` in ONE flat fp32 buffer, written in place by the backward operators, all-reduced
    bucket by bucket WHILE the backward pass is still running (reference: Lightning DDP,
    `ddp_find_unused_parameters_false`, scripts/trainer.yaml:47 — bucketed all-reduce overlapped with backward).

    bf16 compute mode only (its operators accumulate parameter gradients with atomics into zero-filled memory,
    which is what makes a shared pre-zeroed destination possible).  Usage::

        fg = FlatGradients(model.parameters())         # after the process group is up (or single process)
        with fg.profile():                              # ONE ordinary forward + backward: learns the order in which
            loss = step_forward(); loss.backward()      # the pass completes the parameters and how often each is written
        for batch in data:
            fg.step_begin()                             # zero the buffer (before the FORWARD: the fused loss head
            loss = step_forward(); loss.backward()      #   writes its weight gradient during the forward)
            fg.step_end()                               # current stream waits for the bucket all-reduces
            optimizer.step()

    After profile() every ``p.grad`` is a view of the flat buffer, laid out in the order the backward pass finishes
    the parameters and cut into buckets of about ``bucket_bytes``; the operators report each write
    (functional._after_op), a bucket whose last expected write has happened is all-reduced (mean) on a side stream
    from an event recorded at that point.  The sequence is capturable in a CUDA graph (NCCL collectives included:
    capture with ``capture_error_mode="thread_local"``).  With one rank there is no exchange, but the flat layout
    still removes the per-operator gradient slabs and their zero-fill launches.
    Do not call ``zero_grad(set_to_none=True)`` afterwards: it would drop the views."""

    def __init__(self, params, group: Optional[dist.ProcessGroup] = None, bucket_bytes: int = 16 << 20):
        self.params = [p for p in params if p.requires_grad]
        if not self.params:
            raise ValueError("FlatGradients: no trainable parameters")
        self.group = group
        self.world = dist.get_world_size(group) if (dist.is_available() and dist.is_initialized()) else 1
        self.bucket_bytes = int(bucket_bytes)
        self.flat = None
        self.handles = []
        self.fired = 0
        # called as on_ready(bucket, handle) when a bucket's gradients are final on this rank and — more than one rank — its
        # all-reduce has been issued (``handle``: wait on it from the stream that consumes the bucket; None with one rank)
        self.on_ready = None

    # ------------------------------------------------------------------ profiling pass
    def profile(self):
        import contextlib
        from . import functional as F

        @contextlib.contextmanager
        def cm():
            if self.flat is not None:
                raise RuntimeError("FlatGradients.profile(): already built")
            F._TOUCH_LOG = []
            try:
                yield self
                log = F._TOUCH_LOG
            finally:
                F._TOUCH_LOG = None
            self._build(log)
        return cm()

    def _build(self, log):
        from . import functional as F
        dev = self.params[0].device
        by_key = {F._pkey(p): p for p in self.params}
        last, count = {}, {}
        for i, k in enumerate(log):
            if k in by_key:
                last[k] = i
                count[k] = count.get(k, 0) + 1
        touched = sorted((k for k in by_key if k in last), key=lambda k: last[k])
        untouched = [k for k in by_key if k not in last]  # gradients produced by foreign operators (or unused)
        order = touched + untouched
        self.bucket_of, self.expected, self.bounds = {}, [], []
        off, start, nbytes, exp = 0, 0, 0, 0
        offsets = {}
        sizes = [4 * ((by_key[k].numel() + 3) // 4 * 4) for k in order]
        left_touched = sum(sizes[:len(touched)])
        for i, k in enumerate(order):
            p = by_key[k]
            offsets[k] = off
            n = (p.numel() + 3) // 4 * 4
            off += n
            nbytes += 4 * n
            if i < len(touched):
                left_touched -= 4 * n
            exp += count.get(k, 0)
            self.bucket_of[k] = len(self.bounds)
            # a bucket closes when it is full — unless what is left of the pass would make a runt (a collective of a few
            # hundred KB costs the same latency as one of several MB, and the last one is the exposed one)
            full = nbytes >= self.bucket_bytes and (i >= len(touched) or left_touched == 0 or left_touched >= self.bucket_bytes // 4)
            boundary = full or i + 1 == len(order) or (i + 1 == len(touched) and untouched)
            if boundary:
                self.bounds.append((start, off))
                self.expected.append(exp)
                start, nbytes, exp = off, 0, 0
        self.n_touched_buckets = len({self.bucket_of[k] for k in touched})
        self.flat = torch.zeros(off, dtype=torch.float32, device=dev)
        self.views = [self.flat[a:b] for a, b in self.bounds]
        for k, p in by_key.items():
            g = self.flat[offsets[k]:offsets[k] + p.numel()].view_as(p)
            if k in last:
                F._DST[k] = (g.data_ptr(), self)
            p.grad = g  # untouched parameters: autograd accumulates into the view in place, reduced at step_end
        self.comm = torch.cuda.Stream(device=dev) if (self.flat.is_cuda and self.world > 1) else None
        self.remaining = list(self.expected)

    def close(self):
        """unregister the destinations (the gradients stay where they are)"""
        from . import functional as F
        for k in [k for k, v in F._DST.items() if v[1] is self]:
            del F._DST[k]

    # ------------------------------------------------------------------ per step
    def step_begin(self):
        self.flat.zero_()
        self.remaining = list(self.expected)
        self.handles = []
        self.fired = 0

    def _written(self, key):
        b = self.bucket_of[key]
        self.remaining[b] -= 1
        if self.remaining[b] == 0:
            self._reduce(b)

    def bucket_params(self, b):
        """the parameters whose gradients live in bucket ``b``"""
        from . import functional as F
        return [p for p in self.params if self.bucket_of[F._pkey(p)] == b]

    def _reduce(self, b):
        self.fired |= 1 << b
        if self.world == 1:
            if self.on_ready is not None:
                self.on_ready(b, None)
            return
        view = self.views[b]
        nccl = dist.get_backend(self.group) == "nccl"
        op = dist.ReduceOp.AVG if nccl else dist.ReduceOp.SUM
        if self.comm is None:  # CPU tensors (gloo tests)
            self.handles.append(dist.all_reduce(view, op=op, group=self.group, async_op=True))
        else:
            ev = torch.cuda.Event()
            ev.record(torch.cuda.current_stream(view.device))  # everything the backward pass has enqueued so far
            self.comm.wait_event(ev)
            with torch.cuda.stream(self.comm):
                self.handles.append(dist.all_reduce(view, op=op, group=self.group, async_op=True))
        if self.on_ready is not None:
            self.on_ready(b, self.handles[-1])

    def step_end(self):
        """after loss.backward(): reduce what has not been reduced yet (buckets of parameters no operator reported),
        then make the current stream wait for every bucket."""
        for b in range(len(self.views)):
            if not (self.fired >> b) & 1:
                self._reduce(b)
        if self.world == 1:
            return
        for h in self.handles:
            h.wait()
        self.handles = []
        if self.comm is not None:
            torch.cuda.current_stream(self.flat.device).wait_stream(self.comm)
        if dist.get_backend(self.group) != "nccl":
            self.flat.div_(self.world)


def shard_batch(batch: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Contiguous batch shard of this rank (DistributedSampler-style equal split; the remainder, if any,
    goes to the first ranks)."""
    n = batch.shape[0]
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return batch[start:start + base + (1 if rank < rem else 0)]
