"""Phase timestamps of CTA (0, 0) of the one-tile cross-attention backward (libpio_b200_trace.so: python
perceiver-io-1_b200/build.py --trace).  Usage: PIO_B200_LIB=.../libpio_b200_trace.so python tools/xattn_trace.py [b n m heads]"""
import ctypes, os, sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); DEV = "cuda"
b, n, m, heads = (int(x) for x in sys.argv[1:5]) if len(sys.argv) >= 5 else (64, 64, 512, 4)
dh = 64; e = heads * dh
q = torch.randn(b, n, e, device=DEV).to(torch.bfloat16); kv = torch.randn(b, m, 2 * e, device=DEV).to(torch.bfloat16)
do = torch.randn(b, n, e, device=DEV).to(torch.bfloat16)
o = torch.empty(b, n, e, device=DEV); lse = torch.empty(b, heads, n, device=DEV)
d = N.AttnDesc()
d.q, d.k, d.v, d.pad, d.o, d.lse = q.data_ptr(), kv.data_ptr(), kv.data_ptr() + 2 * e, None, o.data_ptr(), lse.data_ptr()
d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * 2 * e, 2 * e, m * 2 * e, 2 * e, n * e, e
d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
st = N.stream_ptr()
N.check(L.pio_attn_fwd(ctypes.byref(d), 1, st))
dq = torch.empty_like(q); dkv = torch.empty_like(kv); delta = torch.empty(b, heads, n, device=DEV)
flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
for _ in range(3):
    flush.zero_()
    N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do.data_ptr(), dq.data_ptr(), dkv.data_ptr(), dkv.data_ptr() + 2 * e, delta.data_ptr(), st))
torch.cuda.synchronize()
slots = 8 * 65
buf = (ctypes.c_longlong * slots)()
raw = ctypes.CDLL(N.LIB_PATH)
raw.pio_debug_xat_trace(buf, slots)
t = list(buf)
t0 = t[8 * 64]
ghz = 1.9
f = lambda v: f"{(v - t0) / ghz / 1000:7.2f}"
print(f"entry 0.00 us, prologue done {f(t[8 * 64 + 1])}, end {f(t[8 * 64 + 2])}  (us at {ghz} GHz)")
names = ["S ready", "S in regs", "prev prod", "drained", "P stored", "mma: S+1 issued", "mma: P seen", "mma: prod issued"]
tiles = (m + 127) // 128
for j in range(tiles):
    print(f"tile {j}: " + "  ".join(f"{names[k]} {f(t[8 * j + k])}" for k in range(8) if t[8 * j + k]))
