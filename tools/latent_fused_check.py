"""Debug aid for the fused latent-block kernel: run pio_latent_block_fwd fused and unfused on the same input and
compare every region of the saved-for-backward buffer (both paths write the same layout), plus the timing of both.

    python tools/latent_fused_check.py [batch] [n] [layers]
"""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

os.environ.setdefault("PIO_FUSED_LATENT_MIN_ROWS", "0")  # the tools always exercise the fused kernel

import perceiver  # noqa: E402
from perceiver import _native as N  # noqa: E402
from perceiver import functional as F  # noqa: E402
from perceiver.model import self_attention_block  # noqa: E402


def al(x, a=256):
    return (x + a - 1) // a * a


def layout(batch, n, ch, heads):
    r = batch * n
    sa, off = {}, 0
    for name, nbytes in (("stats", r * 2 * 4), ("h0", r * ch * 2), ("qkv", 3 * r * ch * 2), ("o", r * ch * 4), ("o16", r * ch * 2),
                         ("lse", batch * heads * n * 4)):
        sa[name] = (off, nbytes)
        off += al(nbytes)
    sa_total = off + 256
    mlp, off = {}, 0
    for name, nbytes in (("stats", r * 2 * 4), ("h0", r * ch * 2), ("h1", r * ch * 2), ("g", r * ch * 2)):
        mlp[name] = (off, nbytes)
        off += al(nbytes)
    mlp_total = off + 256
    act = al(r * ch * 4)
    xmid = al(sa_total)
    mlp_off = xmid + act
    xnext = mlp_off + al(mlp_total)
    per_layer = xnext + act
    regions = {}
    for k, (o, nb) in sa.items():
        regions["sa." + k] = (o, nb, torch.bfloat16 if k in ("h0", "qkv", "o16") else torch.float32)
    regions["xmid"] = (xmid, r * ch * 4, torch.float32)
    for k, (o, nb) in mlp.items():
        regions["mlp." + k] = (mlp_off + o, nb, torch.float32 if k == "stats" else torch.bfloat16)
    regions["xnext"] = (xnext, r * ch * 4, torch.float32)
    return regions, per_layer


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 4
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    layers = int(sys.argv[3]) if len(sys.argv) > 3 else 2
    ch, heads = 256, 4
    perceiver.set_compute_mode("bf16")
    torch.manual_seed(0)
    dev = torch.device("cuda")
    block = self_attention_block(layers, ch, heads, 0.0).to(dev)
    with torch.no_grad():
        for p in block.parameters():
            if p.dim() == 1:
                p.add_(torch.randn_like(p) * 0.1)
    params = []
    for layer in block:
        params.extend(layer[0].module._params())
        params.extend(layer[1].module._params())
    params = [p.detach().contiguous() for p in params]
    x = torch.randn(batch, n, ch, device=dev)
    L = N.lib()
    attn, mlp, keep = F.LatentBlockFn._structs(params, layers, ch, N.MODE_BF16)
    nbytes = L.pio_latent_block_saved_bytes(layers, batch, n, ch, heads, N.MODE_BF16)
    ws = N.workspace(L.pio_latent_block_ws_bytes(layers, batch, n, ch, heads, N.MODE_BF16), dev)

    def run(fused):
        if fused:
            os.environ.pop("PIO_NO_FUSED_LATENT", None)
        else:
            os.environ["PIO_NO_FUSED_LATENT"] = "1"
        saved = torch.zeros(nbytes, dtype=torch.uint8, device=dev)
        y = torch.zeros_like(x)
        def call():
            N.check(L.pio_latent_block_fwd(x.data_ptr(), y.data_ptr(), attn, mlp, saved.data_ptr(), layers, batch, n, ch, heads, None,
                                           N.MODE_BF16, ws.data_ptr(), ws.numel(), N.stream_ptr(dev)), "pio_latent_block_fwd")
        call()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        for _ in range(3):
            call()
        e0.record()
        for _ in range(10):
            call()
        e1.record()
        e1.synchronize()
        return y, saved, e0.elapsed_time(e1) / 10

    y_u, s_u, t_u = run(False)
    y_f, s_f, t_f = run(True)
    print(f"batch {batch} n {n} layers {layers}: unfused {t_u * 1e3:.1f} us, fused {t_f * 1e3:.1f} us per block "
          f"({t_u * 1e3 / layers:.1f} vs {t_f * 1e3 / layers:.1f} us per layer)")
    regions, per_layer = layout(batch, n, ch, heads)

    def rel(a, b):
        a, b = a.double(), b.double()
        return ((a - b).abs().max() / b.abs().max().clamp_min(1e-12)).item(), bool(torch.isfinite(a).all())

    print("y:", rel(y_f, y_u))
    for l in range(layers):
        for name, (off, nb, dt) in regions.items():
            if (name == "xnext" and l == layers - 1) or name == "sa.o":  # (the fused kernel does not write the fp32 copy of O)
                continue
            a = s_f[l * per_layer + off:l * per_layer + off + nb].view(dt)
            b = s_u[l * per_layer + off:l * per_layer + off + nb].view(dt)
            e, fin = rel(a, b)
            flag = "" if (e < 2e-2 and fin) else "   <<<<<<"
            print(f"layer {l} {name:10s} rel err {e:.3e} finite {fin}{flag}")


if __name__ == "__main__":
    main()
