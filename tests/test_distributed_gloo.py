"""Multi-rank host logic on CPU (gloo, world_size 2): the flat gradient all-reduce leaves every rank
with the mean gradient, which for batch-sharded data equals the single-process gradient of the mean
loss over the concatenated batch."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from perceiver.distributed import average_gradients, flatten_gradients, shard_batch


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)  # same "model" on every rank
    lin = torch.nn.Linear(6, 3)
    extra = torch.nn.Parameter(torch.zeros(4))  # a parameter that receives no gradient on any rank
    torch.manual_seed(1)
    data = torch.randn(8, 6)
    target = torch.randn(8, 3)
    x, y = shard_batch(data, rank, world), shard_batch(target, rank, world)
    loss = ((lin(x) - y) ** 2).mean()
    loss.backward()
    params = list(lin.parameters()) + [extra]
    average_gradients(params)
    torch.save({"flat": flatten_gradients(params), "extra_none": extra.grad is None or bool((extra.grad == 0).all())},
               os.path.join(out_dir, f"rank{rank}.pt"))
    dist.barrier()
    dist.destroy_process_group()


def test_shard_batch_covers_the_batch_without_overlap():
    b = torch.arange(10)
    parts = [shard_batch(b, r, 3) for r in range(3)]
    assert torch.equal(torch.cat(parts), b)
    assert [len(p) for p in parts] == [4, 3, 3]


def test_average_gradients_is_a_noop_without_a_process_group():
    p = torch.nn.Parameter(torch.ones(3))
    p.grad = torch.full((3,), 2.0)
    average_gradients([p])
    assert torch.equal(p.grad, torch.full((3,), 2.0))


@pytest.mark.timeout(120)
def test_two_rank_gloo_gradient_mean_equals_full_batch_gradient(tmp_path):
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(tmp_path / f"rank{r}.pt") for r in range(world)]
    assert torch.equal(got[0]["flat"], got[1]["flat"])  # every rank ends with the same gradient
    assert got[0]["extra_none"]
    # reference: one process, whole batch (equal shard sizes -> mean of shard means = full mean)
    torch.manual_seed(0)
    lin = torch.nn.Linear(6, 3)
    torch.manual_seed(1)
    data, target = torch.randn(8, 6), torch.randn(8, 3)
    ((lin(data) - target) ** 2).mean().backward()
    ref = torch.cat([p.grad.reshape(-1) for p in lin.parameters()] + [torch.zeros(4)])
    torch.testing.assert_close(got[0]["flat"], ref, rtol=1e-5, atol=1e-6)
