"""Phase timeline of the cluster version of the fused latent-block kernel (CTA 0 = rank 0 of cluster 0) from a
-DPIO_LB_TRACE build:

    python perceiver-io-1_b200/build.py --trace
    PIO_B200_LIB=perceiver-io-1_b200/libpio_b200_trace.so python tools/latent_cluster_trace.py [batch] [n] [layers]

who 0 = MMA thread: 1000+i before / 2000+i after its i-th wait (per layer: gather LN1, Q/K/V tiles, P, gather O,
gather LN2, gather GELU); who 1 = thread 0 of the element-wise warps: 1000+j / 2000+j around its j-th wait for the
MMA thread (per layer: QKV, S, PV, out-projection, W1, W2), 4000 / 4001 around the LayerNorm statistics exchange,
5000 / 5001 = thread 0 may overwrite the atoms (peers done with the previous operand) / has sent its part; who 0 tag
3000+i = all four atoms of a gather have arrived (1000+i / 2000+i of a gather are around the wait for the CTA's own atom).  Prints the last layer as one merged timeline (cycles since the layer's first mark)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

os.environ["PIO_LATENT_CLUSTER"] = "1"

import perceiver  # noqa: E402
from perceiver import _native as N  # noqa: E402
from perceiver import functional as F  # noqa: E402
from perceiver.model import self_attention_block  # noqa: E402

MMA_WAITS = ["gather LN1", "QKV tiles", "P", "gather O", "gather LN2", "gather GELU"]
EW_WAITS = ["QKV", "S", "PV", "out-proj", "W1", "W2"]


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    layers = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    ch, heads = 256, 4
    perceiver.set_compute_mode("bf16")
    dev = torch.device("cuda")
    torch.manual_seed(0)
    block = self_attention_block(layers, ch, heads, 0.0).to(dev)
    params = []
    for layer in block:
        params.extend(layer[0].module._params())
        params.extend(layer[1].module._params())
    params = [p.detach().contiguous() for p in params]
    x = torch.randn(batch, n, ch, device=dev)
    L = N.lib()
    L.pio_debug_lbc_trace.restype = ctypes.c_int
    L.pio_debug_lbc_trace.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    attn, mlp, keep = F.LatentBlockFn._structs(params, layers, ch, N.MODE_BF16)
    saved = torch.zeros(L.pio_latent_block_saved_bytes(layers, batch, n, ch, heads, N.MODE_BF16), dtype=torch.uint8, device=dev)
    ws = N.workspace(L.pio_latent_block_ws_bytes(layers, batch, n, ch, heads, N.MODE_BF16), dev)
    y = torch.zeros_like(x)

    def call():
        N.check(L.pio_latent_block_fwd(x.data_ptr(), y.data_ptr(), attn, mlp, saved.data_ptr(), layers, batch, n, ch, heads, None,
                                       N.MODE_BF16, ws.data_ptr(), ws.numel(), N.stream_ptr(dev)), "pio_latent_block_fwd")
    for _ in range(3):
        call()
    torch.cuda.synchronize()
    L.pio_debug_lbc_trace(None, 0, 0, 1)
    call()
    torch.cuda.synchronize()
    MAXP = 8192
    ev = []
    for who in (0, 1):
        buf = (ctypes.c_longlong * (2 * MAXP))()
        cnt = L.pio_debug_lbc_trace(buf, MAXP, who, 0)
        ev += [(buf[MAXP + i], who, buf[i]) for i in range(cnt)]
    ev.sort()
    t0, tend = ev[0][0], ev[-1][0]
    print(f"kernel span (CTA 0): {tend - t0} cycles for {layers} layers = {(tend - t0) / layers:.0f} per layer")
    # last layer: from the MMA thread's wait number 6 (layers - 1)
    first = 6 * (layers - 1)
    start = next(t for t, who, tag in ev if who == 0 and tag == 1000 + first)
    # the element-wise side enters the layer a little earlier (its LayerNorm): take its last W2 wake-up before `start`
    ew_prev = [t for t, who, tag in ev if who == 1 and tag == 2000 + first - 1]
    if ew_prev:
        start = min(start, ew_prev[0])
    prev = {0: start, 1: start}
    print(f"{'cycle':>7s}  {'who':4s} {'event':34s} {'since own previous':>18s}")
    for t, who, tag in ev:
        if t < start:
            continue
        kind, idx = tag // 1000, tag % 1000
        if kind in (1, 2):
            names = MMA_WAITS if who == 0 else EW_WAITS
            name = ("wait  " if kind == 1 else "woke  ") + names[idx % 6]
        elif kind == 3:
            name = "all four atoms of " + MMA_WAITS[idx % 6]
        else:
            name = {4000: "stats sent, waiting", 4001: "stats complete", 5000: "atom free, sending", 5001: "sent"}.get(tag, str(tag))
        print(f"{t - start:7d}  {'MMA' if who == 0 else 'EW0':4s} {name:34s} {t - prev[who]:18d}")
        prev[who] = t


if __name__ == "__main__":
    main()
