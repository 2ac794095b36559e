"""torchrun worker of tests/test_gpu_train_step.py::test_two_gpu_overlapped_allreduce... (not collected by pytest).
Every rank runs GraphedMLMStep on its shard; rank 0 also computes the full-batch gradients in one process."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402
import torch.distributed as dist  # noqa: E402


def main():
    import perceiver
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    os.environ.setdefault("NCCL_DEBUG", "WARN")
    dist.init_process_group("nccl", device_id=dev)
    perceiver.set_compute_mode("bf16")
    vocab, seq, c = 1003, 128, 64

    def build():
        torch.manual_seed(0)
        enc = PerceiverEncoder(TextInputAdapter(vocab, seq, c), (32, c), 3, 2, 2, 2)
        dec = PerceiverDecoder(TextOutputAdapter(vocab, seq, c), (32, c), 2)
        return PerceiverMLM(enc, dec, TextMasking(vocab, 1, 2, 3)).to(dev)

    g = torch.Generator().manual_seed(7)
    per = 8
    ids_all = torch.randint(3, vocab, (world * per, seq), generator=g)
    pad_all = torch.zeros(world * per, seq, dtype=torch.bool)
    labels_all = torch.full_like(ids_all, -100)
    sel = torch.rand(ids_all.shape, generator=g) < 0.15
    labels_all[sel] = ids_all[sel]
    # every rank must label the same NUMBER of positions for "mean of the rank means" to equal the full-batch mean
    for r in range(world):
        blk = labels_all[r * per:(r + 1) * per]
        n = int((blk != -100).sum())
        want = 120
        flat = blk.reshape(-1)
        if n > want:
            idx = (flat != -100).nonzero().flatten()[want:]
            flat[idx] = -100
        elif n < want:
            idx = (flat == -100).nonzero().flatten()[:want - n]
            flat[idx] = ids_all[r * per:(r + 1) * per].reshape(-1)[idx]
    sl = slice(rank * per, (rank + 1) * per)
    model = build()
    opt = perceiver.FusedAdamW(model.parameters(), lr=0.0, weight_decay=0.0)
    ids, pad, labels = ids_all[sl].to(dev), pad_all[sl].to(dev), labels_all[sl].to(dev)

    class FixedMasking(torch.nn.Module):  # the batch brings its labels: keep the masking out of the comparison
        def forward(self, x, pad_mask, generator=None):
            return x, labels.clone()
    model.masking = FixedMasking()
    step = perceiver.GraphedMLMStep(model, opt, ids, pad, bucket_bytes=64 << 10)
    assert step.fg is not None and len(step.fg.views) >= 3, "expected several buckets"
    loss = step(ids, pad)
    torch.cuda.synchronize()
    got = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    # all ranks hold the same averaged gradients
    for k, v in got.items():
        ref = v.clone()
        dist.broadcast(ref, 0)
        assert torch.equal(ref, v), f"rank {rank}: {k} differs from rank 0 after the all-reduce"
    if rank == 0:
        perceiver.functional._DST.clear()
        full = build()
        lossf, _ = full(ids_all.to(dev), pad_all.to(dev), masking=False, return_loss=True, labels=labels_all.to(dev))
        lossf.backward()
        torch.cuda.synchronize()
        num = den = 0.0
        for k, p in full.named_parameters():
            num += ((got[k].double() - p.grad.double()) ** 2).sum().item()
            den += (p.grad.double() ** 2).sum().item()
        rel = (num / den) ** 0.5
        assert rel < 5e-3, f"averaged 2-rank gradient vs full-batch gradient: rel L2 {rel}"
        print(f"DDP_WORKER_OK rel={rel:.3e} buckets={len(step.fg.views)} launches_per_replay={step.launches_per_replay}", flush=True)
    # leave like bench.py: release the graph (it holds NCCL kernels) before the communicator, skip destroy_process_group
    torch.cuda.synchronize()
    dist.barrier()
    step.graph.reset()
    torch.cuda.synchronize()
    sys.stdout.flush()
    os._exit(0)


if __name__ == "__main__":
    main()
