"""Decoder-shaped products of the cfg2 step (32768 x 256 x 256): fp32 out + bias + residual, and bf16 out + GELU (for ncu)."""
import sys, ctypes, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); dev = torch.device("cuda"); st = N.stream_ptr(dev)
m, n, k = 32768, 256, 256
x = torch.randn(m, k, device=dev).to(torch.bfloat16); w = (torch.randn(n, k, device=dev) * 0.05).to(torch.bfloat16)
bias = torch.zeros(n, device=dev); res = torch.randn(m, n, device=dev); y = torch.empty(m, n, device=dev)
y16 = torch.empty(m, n, device=dev, dtype=torch.bfloat16); aux = torch.empty(m, n, device=dev, dtype=torch.bfloat16)
ws = torch.empty(1 << 20, dtype=torch.uint8, device=dev)
def gemm(c32=None, c16=None, residual=None, act=0, aux16=None):
    d = N.GemmBf16Desc()
    d.a, d.b, d.m, d.n, d.k, d.lda, d.ldb, d.a_kmajor, d.b_kmajor = x.data_ptr(), w.data_ptr(), m, n, k, k, k, 1, 1
    d.bias = bias.data_ptr()
    if c32 is not None: d.c32, d.ldc32 = c32.data_ptr(), n
    if c16 is not None: d.c16, d.ldc16 = c16.data_ptr(), n
    if residual is not None: d.residual, d.ldr = residual.data_ptr(), n
    if aux16 is not None: d.aux16, d.ldaux = aux16.data_ptr(), n
    d.act = act
    N.check(L.pio_gemm_bf16(ctypes.byref(d), ws.data_ptr(), ws.numel(), st), "gemm")
import time
for _ in range(3):
    gemm(c32=y, residual=res)
    gemm(c16=y16, act=1, aux16=aux)
torch.cuda.synchronize()
for name, fn in (("f32+res", lambda: gemm(c32=y, residual=res)), ("bf16+gelu", lambda: gemm(c16=y16, act=1, aux16=aux))):
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(20): fn()
    e1.record(); e1.synchronize()
    print(name, f"{e0.elapsed_time(e1) / 20 * 1e3:.1f} us")
