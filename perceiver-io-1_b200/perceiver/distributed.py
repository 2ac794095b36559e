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


def average_gradients(params: Iterable[torch.nn.Parameter], group: Optional[dist.ProcessGroup] = None) -> None:
    """In-place mean of every parameter gradient across the ranks of ``group`` with ONE flat all-reduce
    (the gradient set of the MLM config is 45 MB: bandwidth is irrelevant on NVLink, launch count is not).
    Equivalent to what DistributedDataParallel leaves in ``.grad``; a no-op without an initialised
    process group or with a single rank."""
    params = [p for p in params if p.requires_grad]
    if not params or not dist.is_available() or not dist.is_initialized():
        return
    world = dist.get_world_size(group)
    if world == 1:
        return
    flat = flatten_gradients(params)
    if flat.is_cuda and dist.get_backend(group) == "nccl":
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
