"""Debug aid for the fused backward chain of the latent block (pio_latent_bwd_cl.cu): forward + backward of a
self_attention_block with the fused backward on and off, same fused forward; prints the error of dx and of every
parameter gradient, and the time of both backward passes.

    python tools/latent_bwd_check.py [batch] [n] [layers]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

import perceiver  # noqa: E402
from perceiver.model import self_attention_block  # noqa: E402


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    layers = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    perceiver.set_compute_mode("bf16")
    torch.manual_seed(0)
    dev = torch.device("cuda")
    block = self_attention_block(layers, 256, 4, 0.0).to(dev)
    with torch.no_grad():
        for p in block.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    x = torch.randn(batch, n, 256, device=dev)
    dy = torch.randn(batch, n, 256, device=dev)

    def run(fused_bwd):
        os.environ["PIO_LATENT_BWD_CLUSTER"] = "1" if fused_bwd else "0"
        block.zero_grad(set_to_none=True)
        xin = x.clone().requires_grad_(True)
        y = block(xin)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        y.backward(dy)
        e1.record()
        torch.cuda.synchronize()
        return xin.grad.detach().clone(), {k: p.grad.detach().clone() for k, p in block.named_parameters()}, e0.elapsed_time(e1)

    for _ in range(2):
        dx_u, g_u, t_u = run(False)
    for _ in range(2):
        dx_f, g_f, t_f = run(True)
    print(f"batch {batch} n {n} layers {layers}: backward unfused {t_u * 1e3:.0f} us, fused chain {t_f * 1e3:.0f} us (eager, incl. launch overhead)")

    def rel(a, b):
        a, b = a.double(), b.double()
        return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item(), bool(torch.isfinite(a).all())

    e, fin = rel(dx_f, dx_u)
    print(f"dx rel err {e:.3e} finite {fin}")
    for k in g_u:
        e, fin = rel(g_f[k], g_u[k])
        flag = "" if (e < 2e-2 and fin) else "   <<<<<<"
        print(f"{k:40s} rel err {e:.3e} finite {fin}{flag}")


if __name__ == "__main__":
    main()
