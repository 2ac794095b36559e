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
    ident = tuple(id(p) for p in params)
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
                    table=torch.empty(len(grads), 2, dtype=torch.int64, device=dev),
                    staging=torch.empty(len(grads), 2, dtype=torch.int64).pin_memory(), ptrs=None)
        _FLAT_PLANS[ident] = plan
    ptrs = tuple(g.data_ptr() for g in grads)
    if ptrs != plan["ptrs"]:  # eager training hands out new gradient buffers every step; a replayed graph does not
        plan["staging"].copy_(torch.from_numpy(np.asarray(list(zip(ptrs, plan["offs"])), dtype=np.int64)))
        plan["table"].copy_(plan["staging"], non_blocking=True)
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


def shard_batch(batch: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """Contiguous batch shard of this rank (DistributedSampler-style equal split; the remainder, if any,
    goes to the first ranks)."""
    n = batch.shape[0]
    base, rem = divmod(n, world)
    start = rank * base + min(rank, rem)
    return batch[start:start + base + (1 if rank < rem else 0)]
