"""The three flavours of the step's dominant kernel — gemm_tc_kernel at the latent-layer products 4096 x 256 x 256:
forward X W^T (+bias, bf16 out), dX = dY W (fp32 out), dW = dY^T X (k = 4096, fp32 atomics) — twice each, for
`ncu --set full` (PIO_NO_PDL=1).  roofline.traffic of bench.py comes from this capture (profiles/r02_dominant_kernel.json)."""
import ctypes
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
for p in (ROOT, os.path.join(ROOT, "perceiver-io-1_b200")):
    sys.path.insert(0, p)
import torch  # noqa: E402

from perceiver import _native as N  # noqa: E402

L = N.lib()
dev = torch.device("cuda")
st = N.stream_ptr(dev)
rows, c = 4096, 256
x = torch.randn(rows, c, device=dev).to(torch.bfloat16)
w = (torch.randn(c, c, device=dev) * 0.05).to(torch.bfloat16)
dy = torch.randn(rows, c, device=dev).to(torch.bfloat16)
bias = torch.zeros(c, device=dev)
ws = torch.empty(16 << 20, dtype=torch.uint8, device=dev)
y16 = torch.empty(rows, c, device=dev, dtype=torch.bfloat16)
dx = torch.empty(rows, c, device=dev)
dw = torch.zeros(c, c, device=dev)


def gemm(a, b, m, n, k, ak, bk, c16=None, c32=None, acc=0, bias=None):
    d = N.GemmBf16Desc()
    d.a, d.b, d.m, d.n, d.k, d.lda, d.ldb, d.a_kmajor, d.b_kmajor = a.data_ptr(), b.data_ptr(), m, n, k, a.shape[1], b.shape[1], ak, bk
    if c16 is not None:
        d.c16, d.ldc16 = c16.data_ptr(), n
    if c32 is not None:
        d.c32, d.ldc32 = c32.data_ptr(), n
    d.accumulate = acc
    if bias is not None:
        d.bias = bias.data_ptr()
    N.check(L.pio_gemm_bf16(ctypes.byref(d), ws.data_ptr(), ws.numel(), st), "gemm")


for _ in range(2):
    gemm(x, w, rows, c, c, 1, 1, c16=y16, bias=bias)   # forward
    gemm(dy, w, rows, c, c, 1, 0, c32=dx)               # dX
    gemm(dy, x, c, c, rows, 0, 0, c32=dw, acc=2)        # dW
torch.cuda.synchronize()
print("ok")
