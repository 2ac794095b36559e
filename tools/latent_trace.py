"""Phase timeline of the fused latent-block kernel (CTA 0) from a -DPIO_LB_TRACE build:

    python perceiver-io-1_b200/build.py --trace
    PIO_B200_LIB=perceiver-io-1_b200/libpio_b200_trace.so python tools/latent_trace.py [batch] [n] [layers]

who 0 = MMA thread: 1000+i before / 2000+i after its i-th wait for the element-wise warps;
who 1 = element-wise warp 0: 1000+i before / 2000+i after its i-th wait for the MMA thread.
Prints, per hand-off of the LAST traced layer, how long the element-wise phase and the MMA phase took (cycles)."""
import ctypes
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
    L.pio_debug_lb_trace.restype = ctypes.c_int
    L.pio_debug_lb_trace.argtypes = [ctypes.c_void_p, ctypes.c_int, ctypes.c_int, ctypes.c_int]
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
    L.pio_debug_lb_trace(None, 0, 0, 1)
    call()
    torch.cuda.synchronize()
    MAXP = 8192
    tr = {}
    for who in (0, 1):
        buf = (ctypes.c_longlong * (2 * MAXP))()
        cnt = L.pio_debug_lb_trace(buf, MAXP, who, 0)
        tr[who] = [(buf[i], buf[MAXP + i]) for i in range(cnt)]
    mma = {tag: t for tag, t in tr[0]}
    ew = {tag: t for tag, t in tr[1]}
    per_layer = 15
    names = ["LN1 | QKV0", "drainQKV0 | S0", "softmax0 | PV0", "drainO0 | outproj0+QKV1", "drainQKV1 | S1", "softmax1 | PV1",
             "drainO1 | outproj1+QKV2", "drainQKV2 | S2", "softmax2 | PV2", "drainO2 | outproj2+QKV3", "drainQKV3 | S3",
             "softmax3 | PV3", "drainO3 | outproj3", "LN2 | W1", "GELU | W2"]
    t0 = min(t for _, t in tr[0] + tr[1])
    tend = max(t for _, t in tr[0] + tr[1])
    print(f"kernel span (CTA 0): {tend - t0} cycles for {layers} layers = {(tend - t0) / layers:.0f} per layer")
    tot_ew = tot_mma = 0
    l = layers - 1
    print(f"{'hand-off (element-wise | MMA)':36s} {'EW busy':>8s} {'MMA side':>9s}   (MMA thread: woke->next wait)")
    for i in range(per_layer):
        k = l * per_layer + i
        # EW phase i: from EW's return of wait k-1 (or start) to MMA thread's wake from wait_ew k
        ew_start = ew.get(2000 + k - 1, t0) if k > 0 else t0
        mma_wake = mma.get(2000 + k)
        ew_back = ew.get(2000 + k)
        mma_next = mma.get(1000 + k + 1)
        if mma_wake is None or ew_back is None:
            continue
        d_ew = mma_wake - ew_start
        d_mma = ew_back - mma_wake
        tot_ew += d_ew
        tot_mma += d_mma
        issue = (mma_next - mma_wake) if mma_next else -1
        print(f"{names[i]:36s} {d_ew:8d} {d_mma:9d}   {issue}")
    print(f"layer {l}: element-wise {tot_ew}, MMA side {tot_mma} cycles")
    # fine-grained marks (tags >= 3000) of the last layer, relative to the first of them
    for who in (0, 1):
        marks = [(tag, t) for tag, t in tr[who] if tag >= 3000]
        if not marks:
            continue
        per = len(marks) // layers
        last = marks[-per:] if per else marks
        print(f"who {who} marks of the last layer (tag: cycles since previous mark):")
        prev = last[0][1]
        print("  " + "  ".join(f"{tag}:{t - prev if i else 0}" for i, (tag, t) in enumerate(last) if not (prev := t if i == 0 else prev) or True))
        line, prev = [], None
        for tag, t in last:
            line.append(f"{tag}:+{0 if prev is None else t - prev}")
            prev = t
        print("  " + " ".join(line))


if __name__ == "__main__":
    main()
