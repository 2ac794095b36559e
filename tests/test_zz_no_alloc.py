"""pio_layernorm_bwd_noalloc: the LayerNorm backward for callers that cannot allocate.

Who needs it: code that runs while a CUDA stream is being captured or inside a hook where cudaMalloc
(and torch's caching allocator, which may fall through to cudaMalloc) must not be called, plugin
hosts that give a kernel library nothing but input and output buffers, and embedders whose device
memory is fixed after start-up.  Such callers cannot hand pio_layernorm_bwd its workspace, so this
entry point keeps the dgamma / dbeta partial sums in a static device array inside the library.

The test calls it through the C ABI twice in a row, with different shapes, so the second call finds
the first call's leftovers in the shared array, and checks every output against the ordinary entry
point (which gets a torch workspace).  Both run the same kernels on the same inputs; only the number
of row blocks may differ (the built-in array caps it at 128), which changes the summation order of
dgamma / dbeta, hence the fp32-parity tolerance of the other GPU tests rather than bit equality.
This file sorts last on purpose: a failure here cannot hide another test under `-x`.
"""
import pytest
import torch

from gpu_util import FP32_TOL, assert_rel

pytestmark = pytest.mark.gpu
DEV = "cuda"


def _both(N, rows, c, seed, with_residual):
    L = N.lib()
    g = torch.Generator().manual_seed(seed)
    x = (torch.randn(rows, c, generator=g) * 2 + 0.3).to(DEV)
    w = torch.randn(c, generator=g).to(DEV)
    b = torch.randn(c, generator=g).to(DEV)
    dy = torch.randn(rows, c, generator=g).to(DEV)
    dres = torch.randn(rows, c, generator=g).to(DEV) if with_residual else None
    st = N.stream_ptr()
    y = torch.empty_like(x)
    stats = torch.empty(rows, 2, device=DEV)
    N.check(L.pio_layernorm_fwd(x.data_ptr(), w.data_ptr(), b.data_ptr(), y.data_ptr(), stats.data_ptr(), rows, c, 1e-5, st))

    # reference: the ordinary entry point with a caller-provided workspace
    dx0, dw0, db0 = torch.empty_like(x), torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    ws = torch.empty(L.pio_layernorm_bwd_ws_bytes(rows, c), dtype=torch.uint8, device=DEV)
    N.check(L.pio_layernorm_bwd(dy.data_ptr(), x.data_ptr(), stats.data_ptr(), w.data_ptr(), N.ptr(dres), dx0.data_ptr(),
                                dw0.data_ptr(), db0.data_ptr(), 0, rows, c, ws.data_ptr(), ws.numel(), st), "pio_layernorm_bwd")

    # under test: outputs allocated up front, then no allocation of any kind around the call
    dx1, dw1, db1 = torch.empty_like(x), torch.empty(c, device=DEV), torch.empty(c, device=DEV)
    torch.cuda.synchronize()
    before = torch.cuda.memory_allocated()
    N.check(L.pio_layernorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), stats.data_ptr(), w.data_ptr(), N.ptr(dres),
                                        dx1.data_ptr(), dw1.data_ptr(), db1.data_ptr(), 0, rows, c, st),
            "pio_layernorm_bwd_noalloc")
    torch.cuda.synchronize()
    assert torch.cuda.memory_allocated() == before
    assert_rel(dx1, dx0, FP32_TOL, f"dx rows={rows} c={c}")
    assert_rel(dw1, dw0, FP32_TOL, f"dgamma rows={rows} c={c}")
    assert_rel(db1, db0, FP32_TOL, f"dbeta rows={rows} c={c}")
    return dw1, db1, (dy, x, stats, w, dres)


def test_noalloc_layernorm_bwd_matches_workspace_entry_point_twice_in_a_row():
    from perceiver import _native as N
    L = N.lib()
    # first call: wide rows, many row blocks -> fills most of the built-in array
    _both(N, rows=5000, c=1024, seed=1, with_residual=True)
    # second call right after: other shape, fewer blocks -> must not pick up the first call's leftovers
    dw, db, (dy, x, stats, w, dres) = _both(N, rows=37, c=131, seed=2, with_residual=False)
    # accumulate = 1 adds onto the existing parameter gradients
    N.check(L.pio_layernorm_bwd_noalloc(dy.data_ptr(), x.data_ptr(), stats.data_ptr(), w.data_ptr(), None, None,
                                        dw.data_ptr(), db.data_ptr(), 1, 37, 131, N.stream_ptr()))
    dw2, db2, _ = _both(N, rows=37, c=131, seed=2, with_residual=False)
    assert_rel(dw, 2 * dw2, FP32_TOL, "accumulated dgamma")
    assert_rel(db, 2 * db2, FP32_TOL, "accumulated dbeta")


def test_noalloc_rejects_sizes_beyond_the_built_in_scratch():
    from perceiver import _native as N
    L = N.lib()
    t = torch.zeros(4, 2048, device=DEV)
    v = torch.zeros(2048, device=DEV)
    rc = L.pio_layernorm_bwd_noalloc(t.data_ptr(), t.data_ptr(), t.data_ptr(), v.data_ptr(), None, t.data_ptr(), v.data_ptr(),
                                     v.data_ptr(), 0, 4, 2048, N.stream_ptr())
    assert rc == 1
    msg = L.pio_last_error().decode()
    assert "1024" in msg and "2048" in msg
