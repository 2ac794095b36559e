"""Phase timeline of the fused backward chain of the latent block (pio_latent_bwd_cl.cu; CTA 0 = rank 0 of cluster 0)
from a -DPIO_LB_TRACE build:

    python perceiver-io-1_b200/build.py --trace
    PIO_B200_LIB=perceiver-io-1_b200/libpio_b200_trace.so python tools/latent_bwd_trace.py [batch] [n] [layers]

who 0 = MMA thread: 1000+i / 2000+i around its i-th wait, 3000+i = all four atoms of a gather have arrived (per layer:
gather g, gather dH, gather g_mid, dO tile, P / dS tiles, dQ|dK|dV tiles, reduce-scatter); who 1 = thread 0 of the
element-wise warps: 1000+j / 2000+j around its j-th wait for the MMA thread (per layer: dG, dN2, dO, S+dP, dQ/dK/dV,
dN1 partial), 4000 / 4001 around a row-sum exchange, 5000 = its part of a gather is written, 6000 / 6001 around the wait
for the reduce-scatter's incoming slices.  Prints the middle layer as one merged timeline."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

os.environ["PIO_LATENT_BWD_CLUSTER"] = "1"

import perceiver  # noqa: E402
from perceiver import _native as N  # noqa: E402
from perceiver.model import self_attention_block  # noqa: E402

MMA_WAITS = ["gather g", "gather dH", "gather g_mid", "dO tile", "P/dS tiles", "dQKV tiles", "reduce-scatter"]
EW_WAITS = ["dG", "dN2", "dO", "S+dP", "dQ/dK/dV", "dN1 partial"]


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 64
    n = int(sys.argv[2]) if len(sys.argv) > 2 else 64
    layers = int(sys.argv[3]) if len(sys.argv) > 3 else 6
    perceiver.set_compute_mode("bf16")
    dev = torch.device("cuda")
    torch.manual_seed(0)
    block = self_attention_block(layers, 256, 4, 0.0).to(dev)
    x = torch.randn(batch, n, 256, device=dev)
    dy = torch.randn(batch, n, 256, device=dev)
    L = N.lib()
    L.pio_debug_lbb_trace.restype = ctypes.c_int
    L.pio_debug_lbb_trace.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
    for it in range(3):
        block.zero_grad(set_to_none=True)
        xin = x.clone().requires_grad_(True)
        y = block(xin)
        torch.cuda.synchronize()
        L.pio_debug_lbb_trace(None, 0, 0, 1)
        y.backward(dy)
        torch.cuda.synchronize()
    L.pio_debug_lbb_cta_times.restype = ctypes.c_int
    L.pio_debug_lbb_cta_times.argtypes = [ctypes.c_void_p, ctypes.c_int]
    tb = (ctypes.c_ulonglong * (4 * 1024))()
    L.pio_debug_lbb_cta_times(tb, 1024)
    ctas = 4 * ((batch * n + 127) // 128)
    st = [[tb[i * 1024 + c] for c in range(ctas)] for i in range(4)]
    base = min(st[0])
    names = ["kernel entry", "set-up done", "work done", "exit"]
    print(f"per-CTA wall clock (ns since the first CTA entered), {ctas} CTAs:")
    for i in range(4):
        v = [x - base for x in st[i]]
        print(f"  {names[i]:12s} min {min(v):7d}  median {sorted(v)[len(v) // 2]:7d}  max {max(v):7d}")
    MAXP = 8192
    ev = []
    for who in (0, 1):
        buf = (ctypes.c_longlong * (2 * MAXP))()
        cnt = L.pio_debug_lbb_trace(buf, MAXP, who, 0)
        ev += [(buf[MAXP + i], who, buf[i]) for i in range(cnt)]
    ev.sort()
    t0, tend = ev[0][0], ev[-1][0]
    print(f"kernel span (CTA 0): {tend - t0} cycles for {layers} layers = {(tend - t0) / layers:.0f} per layer")
    it = layers // 2  # the it-th processed layer
    start = next(t for t, who, tag in ev if who == 0 and tag == 1000 + 7 * it)
    stop = next((t for t, who, tag in ev if who == 0 and tag == 1000 + 7 * (it + 1)), tend)
    prev = {0: start, 1: start}
    print(f"{'cycle':>7s}  {'who':4s} {'event':34s} {'since own previous':>18s}")
    for t, who, tag in ev:
        if t < start or t > stop:
            continue
        kind, idx = tag // 1000, tag % 1000
        if kind in (1, 2):
            names, per = (MMA_WAITS, 7) if who == 0 else (EW_WAITS, 6)
            name = ("wait  " if kind == 1 else "woke  ") + names[idx % per]
        elif kind == 3:
            name = "all four atoms of " + MMA_WAITS[idx % 7]
        else:
            name = {4000: "row sums sent, waiting", 4001: "row sums complete", 5000: "part written, sending", 6000: "waiting for slices",
                    6001: "slices received"}.get(tag, str(tag))
        print(f"{t - start:7d}  {'MMA' if who == 0 else 'EW0':4s} {name:34s} {t - prev[who]:18d}")
        prev[who] = t


if __name__ == "__main__":
    main()
