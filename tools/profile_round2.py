"""One pass over the kernels added in the second half of round 2, for ncu:

    ncu --set full --clock-control none --import-source on \
        -k regex:'latent_block_(fwd|bwd)_cluster_kernel|dw_batched_kernel|ln_param_grads_kernel|logits_persistent_kernel|ce_chunk_reg' \
        -o gpurun_out/r02b python tools/profile_round2.py

Runs, at the benchmarked cfg2 shapes: forward + backward of one latent block (64 samples x 64 latents x 256 channels, 6
layers: the forward cluster kernel, the backward chain kernel, the batched weight-gradient and LayerNorm-gradient
kernels) and the fused loss head on 4915 labelled rows of 32768 (persistent logits kernel, cross-entropy chunk kernel)."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import perceiver  # noqa: E402
from perceiver import functional as F  # noqa: E402
from perceiver.model import self_attention_block  # noqa: E402


def main():
    perceiver.set_compute_mode("bf16")
    torch.manual_seed(0)
    dev = torch.device("cuda")
    block = self_attention_block(6, 256, 4, 0.0).to(dev)
    x = torch.randn(64, 64, 256, device=dev, requires_grad=True)
    dy = torch.randn(64, 64, 256, device=dev)
    for _ in range(2):
        block.zero_grad(set_to_none=True)
        y = block(x)
        y.backward(dy)
    torch.cuda.synchronize()
    # the loss head: 32768 rows, ~15 % labelled
    rows, c, vocab = 32768, 256, 10003
    h = torch.randn(rows, c, device=dev, requires_grad=True)
    w = (torch.randn(vocab, c, device=dev) * 0.05).requires_grad_(True)
    b = torch.zeros(vocab, device=dev, requires_grad=True)
    labels = torch.full((rows,), -100, device=dev, dtype=torch.int64)
    pick = torch.randperm(rows, device=dev)[:4915]
    labels[pick] = torch.randint(0, vocab, (4915,), device=dev)
    for _ in range(2):
        loss = F.linear_cross_entropy(h, w, b, labels)
        loss.backward()
    torch.cuda.synchronize()
    print("done", float(loss))


if __name__ == "__main__":
    main()
