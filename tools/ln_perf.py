"""LayerNorm backward timing at the latent / decoder shapes (atomic accumulate mode, as the bf16 pipelines use it)."""
import sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); dev = torch.device("cuda"); st = N.stream_ptr(dev)
for rows, c in [(4096, 256), (32768, 256), (8192, 512), (32768, 64), (4096, 128)]:
    x = torch.randn(rows, c, device=dev); dy = torch.randn_like(x); dres = torch.randn_like(x); dx = torch.empty_like(x)
    w = torch.randn(c, device=dev); stats = torch.stack([x.mean(1), 1 / (x.var(1, unbiased=False) + 1e-5).sqrt()], 1).contiguous()
    dw = torch.zeros(c, device=dev); db = torch.zeros(c, device=dev)
    ws = torch.empty(L.pio_layernorm_bwd_ws_bytes(rows, c), dtype=torch.uint8, device=dev)
    for acc in (2, 0):
        def run():
            N.check(L.pio_layernorm_bwd(dy.data_ptr(), x.data_ptr(), stats.data_ptr(), w.data_ptr(), dres.data_ptr(), dx.data_ptr(),
                                        dw.data_ptr(), db.data_ptr(), acc, rows, c, ws.data_ptr(), ws.numel(), N.stream_ptr(dev)), "ln")
        for _ in range(5): run()
        g = torch.cuda.CUDAGraph()
        with torch.cuda.graph(g):
            for _ in range(50): run()
        g.replay(); torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); g.replay(); e1.record(); e1.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / 50
        print(f"rows {rows:6d} c {c:4d} accumulate {acc}: {us:6.2f} us/call  ({(4 * rows * c * 4) / us / 1e6:.2f} TB/s algorithmic)")
