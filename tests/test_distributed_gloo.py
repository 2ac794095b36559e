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


# ---------------------------------------------------------------------------------------------------------------
# FlatGradients: gradients written in place into one flat buffer, bucketed all-reduce started from inside the
# backward pass.  The CUDA operators are replaced by a CPU stand-in that asks for its gradient buffers exactly the
# way they do (functional._grads_like / _out) and adds into the address it is given.
# ---------------------------------------------------------------------------------------------------------------
import ctypes  # noqa: E402

import numpy as np  # noqa: E402

from perceiver import _native as N  # noqa: E402
from perceiver import functional as F  # noqa: E402
from perceiver.distributed import FlatGradients  # noqa: E402


def _add_at(buf, value):
    """buf: a tensor or functional._Alias (an address): buf += value, like a kernel with accumulate = 2"""
    if isinstance(buf, torch.Tensor):
        buf.add_(value)
    else:
        n = value.numel()
        arr = np.ctypeslib.as_array((ctypes.c_float * n).from_address(buf.data_ptr()))
        arr += value.detach().reshape(-1).numpy()


class _FakeLinear(torch.autograd.Function):
    """y = x w^T with the gradient-buffer protocol of the fused operators"""

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(x, w)
        return x @ w.t()

    @staticmethod
    def backward(ctx, dy):
        x, w = ctx.saved_tensors
        (gw,), acc = F._grads_like([w], N.MODE_BF16, ctx.needs_input_grad[1:2])
        assert acc == 2
        _add_at(gw, dy.t() @ x)
        return F._out(dy @ w, gw)


def _fg_model():
    torch.manual_seed(0)
    w1 = torch.nn.Parameter(torch.randn(5, 6) * 0.3)   # used twice (like the shared layer_n)
    w2 = torch.nn.Parameter(torch.randn(4, 5) * 0.3)
    w3 = torch.nn.Parameter(torch.randn(3, 4) * 0.3)
    foreign = torch.nn.Parameter(torch.randn(3) * 0.1)  # gradient from an ordinary torch op
    unused = torch.nn.Parameter(torch.zeros(7))
    return [w1, w2, w3, foreign, unused]


def _fg_loss(params, x, y):
    w1, w2, w3, foreign, _ = params
    h = _FakeLinear.apply(x, w1)
    h = torch.tanh(_FakeLinear.apply(h @ torch.ones(5, 6), w1))
    h = torch.tanh(_FakeLinear.apply(h, w2))
    out = _FakeLinear.apply(h, w3) + foreign
    return ((out - y) ** 2).mean()


def _fg_worker(rank, world, port, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    F.set_compute_mode("bf16")
    params = _fg_model()
    torch.manual_seed(1)
    data, target = torch.randn(8, 6), torch.randn(8, 3)
    x, y = shard_batch(data, rank, world), shard_batch(target, rank, world)
    fg = FlatGradients(params, bucket_bytes=64)  # tiny buckets: several of them, cut in completion order
    with fg.profile():
        _fg_loss(params, x, y).backward()
    fired_inside = []
    orig = fg._reduce
    fg._reduce = lambda b: (fired_inside.append(b), orig(b))[1]
    # the per-bucket consumer hook (GraphedTrainStep's optimizer-per-bucket): one call per bucket and step, with the handle
    # of the bucket's all-reduce; after waiting on it the bucket holds the sum over the ranks
    ready, sums = [], {}

    def on_ready(b, handle):
        handle.wait()
        ready.append(b)
        sums[b] = fg.views[b].clone()
    fg.on_ready = on_ready
    results = []
    for step in range(2):  # twice: the per-step state resets
        ready.clear()
        fg.step_begin()
        _fg_loss(params, x, y).backward()
        n_in_backward = len(fired_inside)
        fg.step_end()
        results.append(torch.cat([p.grad.reshape(-1) for p in params]).clone())
        fired_inside.clear()
    hook_ok = (sorted(ready) == list(range(len(fg.views))) and
               all(torch.allclose(sums[b] / world, fg.views[b]) for b in range(len(fg.views))) and
               sorted(id(p) for b in range(len(fg.views)) for p in fg.bucket_params(b)) == sorted(id(p) for p in fg.params))
    torch.save({"g0": results[0], "g1": results[1], "buckets": len(fg.views), "fired_in_backward": n_in_backward,
                "views": all(p.grad.data_ptr() >= fg.flat.data_ptr() for p in params), "hook_ok": hook_ok},
               os.path.join(out_dir, f"fg{rank}.pt"))
    fg.close()
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(120)
def test_flat_gradients_two_rank_gloo_bucketed_mean_equals_full_batch_gradient(tmp_path):
    world = 2
    mp.spawn(_fg_worker, args=(world, _free_port(), str(tmp_path)), nprocs=world, join=True)
    got = [torch.load(tmp_path / f"fg{r}.pt") for r in range(world)]
    assert torch.equal(got[0]["g0"], got[1]["g0"]) and torch.equal(got[0]["g0"], got[0]["g1"])
    assert got[0]["views"] and got[0]["buckets"] >= 3
    assert got[0]["hook_ok"] and got[1]["hook_ok"]
    assert got[0]["fired_in_backward"] >= 2  # buckets went out while the backward pass was still running
    # reference: one process, whole batch, plain autograd
    params = _fg_model()
    torch.manual_seed(1)
    data, target = torch.randn(8, 6), torch.randn(8, 3)
    w1, w2, w3, foreign, unused = params
    h = data @ w1.t()
    h = torch.tanh((h @ torch.ones(5, 6)) @ w1.t())
    h = torch.tanh(h @ w2.t())
    (((h @ w3.t() + foreign) - target) ** 2).mean().backward()
    ref = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1) for p in params])
    torch.testing.assert_close(got[0]["g0"], ref, rtol=1e-5, atol=1e-6)


def test_flat_gradients_single_process_matches_autograd():
    F.set_compute_mode("bf16")
    try:
        params = _fg_model()
        torch.manual_seed(2)
        x, y = torch.randn(4, 6), torch.randn(4, 3)
        fg = FlatGradients(params, bucket_bytes=1 << 20)
        with fg.profile():
            _fg_loss(params, x, y).backward()
        fg.step_begin()
        _fg_loss(params, x, y).backward()
        fg.step_end()
        got = [p.grad.clone() for p in params]
        fg.close()
        assert not F._DST
        for p in params:
            p.grad = None
        _fg_loss(params, x, y).backward()
        for g, p in zip(got, params):
            ref = p.grad if p.grad is not None else torch.zeros_like(p)
            torch.testing.assert_close(g, ref, rtol=1e-5, atol=1e-6)
    finally:
        F.set_compute_mode("fp32")
