"""Out-of-bounds guard: the newer multi-kernel entry points run with every output and the workspace carved out of one
buffer with sentinel gaps around them; the sentinels must survive (compute-sanitizer is not available on the pool)."""
import ctypes

import pytest
import torch

from gpu_util import rel_err

pytestmark = pytest.mark.gpu
DEV = "cuda"
GAP = 4096
SENTINEL = 0xA5


class Arena:
    """sub-buffers of one byte buffer, each between two sentinel gaps"""

    def __init__(self, total_bytes):
        self.buf = torch.full((total_bytes,), SENTINEL, dtype=torch.uint8, device=DEV)
        self.off = GAP
        self.used = []

    def take(self, nbytes, dtype=torch.uint8):
        start = (self.off + 255) // 256 * 256
        end = start + nbytes
        assert end + GAP <= self.buf.numel()
        self.used.append((start, end))
        self.off = end + GAP
        return self.buf[start:end].view(dtype)

    def check(self):
        mask = torch.ones(self.buf.numel(), dtype=torch.bool, device=DEV)
        for a, b in self.used:
            mask[a:b] = False
        torch.cuda.synchronize()
        bad = (self.buf[mask] != SENTINEL).nonzero()
        assert bad.numel() == 0, f"{bad.numel()} sentinel bytes overwritten (first at masked index {int(bad[0])})"


@pytest.mark.parametrize("rows,c,classes,p_ignore", [(300, 64, 517, 0.8), (9001, 72, 1001, 0.05), (257, 136, 40, 0.0), (130, 64, 10003, 0.5)])
def test_linear_ce_stays_inside_its_buffers(rows, c, classes, p_ignore):
    from perceiver import _native as N
    L = N.lib()
    torch.manual_seed(0)
    ws_bytes = L.pio_linear_ce_ws_bytes(rows, c, classes, 1)
    sizes = [rows * c * 4, classes * c * 2, classes * 4, rows * 8, 12, rows * c * 4, classes * c * 4, classes * 4, ws_bytes]
    ar = Arena(sum(sizes) + (len(sizes) + 2) * (GAP + 256))
    x = ar.take(rows * c * 4, torch.float32).view(rows, c)
    x.copy_(torch.randn(rows, c, device=DEV))
    w16 = ar.take(classes * c * 2, torch.bfloat16).view(classes, c)
    w16.copy_((torch.randn(classes, c, device=DEV) * 0.1).to(torch.bfloat16))
    b = ar.take(classes * 4, torch.float32)
    b.copy_(torch.randn(classes, device=DEV) * 0.1)
    labels = ar.take(rows * 8, torch.int64)
    y = torch.randint(0, classes, (rows,), device=DEV)
    y[torch.rand(rows, device=DEV) < p_ignore] = -100
    y[0] = 1
    labels.copy_(y)
    out3 = ar.take(12, torch.float32)
    dx = ar.take(rows * c * 4, torch.float32)
    dw = ar.take(classes * c * 4, torch.float32)
    db = ar.take(classes * 4, torch.float32)
    ws = ar.take(ws_bytes)
    N.check(L.pio_linear_ce_fwd_bwd(x.data_ptr(), w16.data_ptr(), b.data_ptr(), labels.data_ptr(), out3.data_ptr(), dx.data_ptr(),
                                    dw.data_ptr(), db.data_ptr(), rows, c, classes, -100, 1, ws.data_ptr(), ws_bytes,
                                    N.stream_ptr(x.device)), "pio_linear_ce_fwd_bwd")
    ar.check()
    assert torch.isfinite(out3).all() and torch.isfinite(dx).all() and torch.isfinite(dw).all() and torch.isfinite(db).all()
    assert int(out3[1].item()) == int((y != -100).sum().item())


def test_adamw_and_flat_copy_stay_inside_their_tensors():
    import numpy as np
    from perceiver import _native as N
    L = N.lib()
    shapes = [(3, 131), (517,), (1,), (16384 + 5,), (40, 72)]
    n_el = [int(np.prod(s)) for s in shapes]
    ar = Arena(sum(n * 4 * 5 for n in n_el) + sum(n_el) * 4 + 3 * 136 * 2 + 40 * 72 * 2 + 64 * GAP + 65536)
    tensors, chunks, flat_rows, off = [], [], [], 0
    keep = []
    for ti, (s, n) in enumerate(zip(shapes, n_el)):
        p, g, m, v = (ar.take(n * 4, torch.float32) for _ in range(4))
        p.copy_(torch.randn(n, device=DEV)); g.copy_(torch.randn(n, device=DEV)); m.zero_(); v.zero_()
        sh, cols, ld = None, 0, 0
        if len(s) == 2:
            cols, ld = s[1], (s[1] + 7) // 8 * 8
            sh = ar.take(s[0] * ld * 2, torch.bfloat16)
            sh.zero_()
        keep.append((p, g, m, v, sh))
        tensors.append((p.data_ptr(), g.data_ptr(), m.data_ptr(), v.data_ptr(), sh.data_ptr() if sh is not None else 0, cols | (ld << 32)))
        flat_rows.append((g.data_ptr(), off))
        for o in range(0, n, 16384):
            chunks.append((ti | (min(16384, n - o) << 32), o))
        off += (n + 3) // 4 * 4
    t_dev = torch.from_numpy(np.asarray(tensors, dtype=np.int64)).to(DEV)
    c_dev = torch.from_numpy(np.asarray(chunks, dtype=np.int64).reshape(-1, 2)).to(DEV)
    f_dev = torch.from_numpy(np.asarray(flat_rows, dtype=np.int64)).to(DEV)
    hyper = torch.tensor([1e-2, 0.9, 0.999, 1e-8, 0.01, 0.0], device=DEV)
    flat = ar.take(off * 4, torch.float32)
    st = N.stream_ptr(torch.device(DEV))
    before = [k[0].clone() for k in keep]
    N.check(L.pio_flat_copy_f32(flat.data_ptr(), f_dev.data_ptr(), c_dev.data_ptr(), len(chunks), 1, 1.0, st), "gather")
    N.check(L.pio_flat_copy_f32(flat.data_ptr(), f_dev.data_ptr(), c_dev.data_ptr(), len(chunks), 0, 0.5, st), "scatter")
    N.check(L.pio_adamw_step(t_dev.data_ptr(), c_dev.data_ptr(), len(chunks), hyper.data_ptr(), st), "adamw")
    N.check(L.pio_adamw_tick(hyper.data_ptr(), st), "tick")
    ar.check()
    assert float(hyper[5].item()) == 1.0
    for (p, g, m, v, sh), s, b4 in zip(keep, shapes, before):
        assert not torch.equal(p, b4) and torch.isfinite(p).all()
        if sh is not None:
            ld = (s[1] + 7) // 8 * 8
            assert torch.equal(sh.view(s[0], ld)[:, :s[1]], p.view(s).to(torch.bfloat16))
            assert not sh.view(s[0], ld)[:, s[1]:].any()


def test_two_threads_on_two_streams_run_backward_operators_concurrently():
    """the side streams / fork-join events of the bf16 backward operators were one set per device, so
    two host threads (or two streams) could overwrite each other's fork record.  They are now keyed by (device, caller
    stream): two threads hammering pio_mlp_bwd (accumulate = 2, side-stream weight gradients) on their own streams must
    each reproduce their single-threaded result."""
    import threading
    import perceiver
    from perceiver import functional as F
    perceiver.set_compute_mode("bf16")
    try:
        torch.manual_seed(0)
        c, rows = 128, 2048

        def make(seed):
            g = torch.Generator(device=DEV).manual_seed(seed)
            x = torch.randn(rows, c, device=DEV, generator=g)
            ps = [torch.randn(s, device=DEV, generator=g) * 0.2 for s in [(c,), (c,), (c, c), (c,), (c, c), (c,)]]
            dy = torch.randn(rows, c, device=DEV, generator=g)
            return x, ps, dy

        def run(x, ps, dy, iters, out):
            st = torch.cuda.Stream()
            with torch.cuda.stream(st):
                for _ in range(iters):
                    xin = x.clone().requires_grad_(True)
                    pin = [p.clone().requires_grad_(True) for p in ps]
                    y = F.MlpFn.apply(xin, *pin, True)
                    y.backward(dy)
                    out.append([xin.grad.clone()] + [p.grad.clone() for p in pin])
            st.synchronize()

        cases = [make(1), make(2)]
        ref = []
        for cs in cases:
            o = []
            run(*cs, 1, o)
            ref.append(o[0])
        outs = [[], []]
        threads = [threading.Thread(target=run, args=(*cases[i], 25, outs[i])) for i in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        torch.cuda.synchronize()
        for i in range(2):
            assert len(outs[i]) == 25
            for got in outs[i]:
                for g, r in zip(got, ref[i]):
                    assert rel_err(g, r) < 2e-3  # (atomics: summation order differs from run to run)
    finally:
        perceiver.set_compute_mode("fp32")
