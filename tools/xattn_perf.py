"""cross-attention backward at the benchmark shape (64 samples, 64 latents, 512 keys, 4 heads of 64): the one-tile kernel
(pio_xattn_tc.cu) against the two general kernels (PIO_NO_XATTN_TC=1), timed alone with CUDA events."""
import ctypes, os, sys, torch
sys.path.insert(0, '/root/repo'); sys.path.insert(0, '/root/repo/perceiver-io-1_b200')
from perceiver import _native as N
L = N.lib(); DEV = "cuda"
def run(b, n, m, heads, reps=20, dh=64):
    e = heads * dh
    q = torch.randn(b, n, e, device=DEV).to(torch.bfloat16); kv = torch.randn(b, m, 2 * e, device=DEV).to(torch.bfloat16)
    do = torch.randn(b, n, e, device=DEV).to(torch.bfloat16)
    o = torch.empty(b, n, e, device=DEV); lse = torch.empty(b, heads, n, device=DEV)
    d = N.AttnDesc()
    d.q, d.k, d.v, d.pad, d.o, d.lse = q.data_ptr(), kv.data_ptr(), kv.data_ptr() + 2 * e, None, o.data_ptr(), lse.data_ptr()
    d.q_bs, d.q_rs, d.k_bs, d.k_rs, d.v_bs, d.v_rs, d.o_bs, d.o_rs = n * e, e, m * 2 * e, 2 * e, m * 2 * e, 2 * e, n * e, e
    d.batch, d.heads, d.n, d.m, d.dh, d.scale = b, heads, n, m, dh, dh ** -0.5
    st = N.stream_ptr()
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=DEV)
    N.check(L.pio_attn_fwd(ctypes.byref(d), 1, st))
    dq = torch.empty_like(q); dkv = torch.empty_like(kv); delta = torch.empty(b, heads, n, device=DEV)
    def t(fn):
        ts = []
        for i in range(reps + 3):
            flush.zero_()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); fn(); e1.record(); e1.synchronize()
            if i >= 3: ts.append(e0.elapsed_time(e1))
        ts.sort()
        return ts[len(ts) // 2]
    res = {}
    for name, env in (("one-tile", None), ("general", "1")):
        if env: os.environ["PIO_NO_XATTN_TC"] = env
        else: os.environ.pop("PIO_NO_XATTN_TC", None)
        res[name] = t(lambda: N.check(L.pio_attn_bwd_bf16(ctypes.byref(d), do.data_ptr(), dq.data_ptr(), dkv.data_ptr(), dkv.data_ptr() + 2 * e, delta.data_ptr(), st)))
    os.environ.pop("PIO_NO_XATTN_TC", None)
    print(f"B{b} N{n} M{m} H{heads}: " + "  ".join(f"{k} {v * 1e3:7.1f} us" for k, v in res.items()))
run(64, 64, 512, 4)
run(64, 64, 512, 4)
run(64, 128, 512, 4)
run(64, 64, 2048, 4)
run(16, 64, 512, 4)
