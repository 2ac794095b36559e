"""The step's three vocabulary-head products at the in-step chunk shape, plus one latent-layer product, once each
(for `ncu --set full`; PIO_NO_PDL=1)."""
import sys, ctypes, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); dev = torch.device("cuda"); st = N.stream_ptr(dev)
rows, c, v = 4352, 256, 10003; vp = 10008
x = torch.randn(rows, c, device=dev).to(torch.bfloat16); w = (torch.randn(v, c, device=dev) * 0.05).to(torch.bfloat16)
bias = torch.zeros(v, device=dev)
dl = torch.zeros(rows, vp, device=dev, dtype=torch.bfloat16); dl[:, :v] = torch.randn(rows, v, device=dev).to(torch.bfloat16)
ws = torch.empty(64 << 20, dtype=torch.uint8, device=dev)
def gemm(a, b, m, n, k, ak, bk, c16=None, ldc16=0, c32=None, acc=0, bias=None):
    d = N.GemmBf16Desc()
    d.a, d.b, d.m, d.n, d.k, d.lda, d.ldb, d.a_kmajor, d.b_kmajor = a.data_ptr(), b.data_ptr(), m, n, k, a.shape[1], b.shape[1], ak, bk
    if c16 is not None: d.c16, d.ldc16 = c16.data_ptr(), ldc16
    if c32 is not None: d.c32, d.ldc32 = c32.data_ptr(), n
    d.accumulate = acc
    if bias is not None: d.bias = bias.data_ptr()
    N.check(L.pio_gemm_bf16(ctypes.byref(d), ws.data_ptr(), ws.numel(), st), "gemm")
lg = torch.empty(rows, vp, device=dev, dtype=torch.bfloat16); dx = torch.empty(rows, c, device=dev); dw = torch.zeros(v, c, device=dev)
xl = torch.randn(4096, 256, device=dev).to(torch.bfloat16); wl = torch.randn(768, 256, device=dev).to(torch.bfloat16)
ql = torch.empty(4096, 768, device=dev, dtype=torch.bfloat16); bl = torch.zeros(768, device=dev)
for _ in range(2):
    gemm(x, w, rows, v, c, 1, 1, c16=lg, ldc16=vp, bias=bias)      # logits chunk
    gemm(dl, w, rows, c, v, 1, 0, c32=dx)                           # dx chunk
    gemm(dl, x, v, c, rows, 0, 0, c32=dw, acc=2)                    # dW chunk (atomic accumulate)
    gemm(xl, wl, 4096, 768, 256, 1, 1, c16=ql, ldc16=768, bias=bl)  # latent QKV projection
torch.cuda.synchronize(); print("ok")
