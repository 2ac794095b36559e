"""one launch of the one-tile cross-attention backward at the benchmark shape (64 x 4 heads, 64 queries, 512 keys), for ncu"""
import ctypes, sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); DEV = "cuda"
b, n, m, heads, dh = 64, 64, 512, 4, 64
e = heads * dh
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
for _ in range(2):
    N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do.data_ptr(), dq.data_ptr(), dkv.data_ptr(), dkv.data_ptr() + 2 * e, delta.data_ptr(), st))
torch.cuda.synchronize()
print("ok")
