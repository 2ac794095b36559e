"""Host-side bookkeeping of perceiver.functional that needs no GPU: when bf16 weight shadows are re-cast, and how
gradient buffers of a parameter used twice in one forward are shared.  The native library is replaced by a recorder."""
import pytest
import torch

import perceiver  # noqa: F401  (puts the package on the path via conftest)
from perceiver import _native as N
from perceiver import functional as F


class _Recorder:
    def __init__(self):
        self.single = 0
        self.multi = []

    def pio_cast_bf16(self, *a):
        self.single += 1
        return 0

    def pio_cast_bf16_multi(self, table, count, stream):
        self.multi.append(count)
        return 0


@pytest.fixture
def fake(monkeypatch):
    rec = _Recorder()
    monkeypatch.setattr(N, "lib", lambda: rec)
    monkeypatch.setattr(N, "stream_ptr", lambda device=None: 0)
    monkeypatch.setattr(torch.cuda, "is_current_stream_capturing", lambda: False)
    F.set_compute_mode("bf16")
    F._forget_pending()
    yield rec
    F.set_compute_mode("fp32")
    F._forget_pending()


class _Toy(torch.nn.Module):
    def __init__(self):
        super().__init__()
        self.a = torch.nn.Parameter(torch.randn(8, 16))
        self.b = torch.nn.Parameter(torch.randn(4, 9))  # pitch 9 -> 16
        self.frozen = torch.nn.Parameter(torch.randn(8, 8), requires_grad=False)

    def forward(self):
        with F.shadow_scope(self):
            return [F._w16(self.a), F._w16(self.b), F._w16(self.a), F._w16(self.frozen)]


def test_training_forwards_recast_once_per_forward_then_in_one_launch(fake):
    m = _Toy()
    s1 = m()
    assert fake.single == 3 and fake.multi == []          # a, b, frozen; the second use of `a` hits the forward's own cast
    assert s1[0] is s1[2] and s1[1].shape == (4, 16)
    s2 = m()
    assert fake.single == 3 and fake.multi == [2]          # a and b in ONE launch, frozen served from the version cache
    assert s2[0] is s1[0] and s2[1] is s1[1]               # refreshed in place (captured graphs keep their addresses)
    with torch.no_grad():
        m.frozen.add_(1.0)                                 # version moves
    m()
    assert fake.single == 4 and fake.multi == [2, 2]


def test_no_scope_means_no_trust_for_trainable_weights(fake):
    m = _Toy()
    F._w16(m.a)
    F._w16(m.a)
    assert fake.single == 2                                # fused optimizers do not move version counters
    F._w16(m.frozen)
    F._w16(m.frozen)
    assert fake.single == 3


def test_eval_after_training_does_not_reuse_the_training_shadow(fake):
    m = _Toy()
    m()
    m()
    n = fake.single
    with torch.no_grad():
        m()                                                # entries were made in training mode: re-cast once ...
        assert fake.single == n + 2
        m()                                                # ... then cached by version
        assert fake.single == n + 2


def test_shadows_written_by_the_optimizer_are_not_cast_again(fake):
    m = _Toy()
    m()
    m()
    assert fake.multi == [2]
    for p in (m.a, m.b):
        p._pio_fresh16 = True                              # what perceiver.FusedAdamW.step() leaves behind
    m()
    assert fake.multi == [2]
    m()                                                    # flags are consumed by one forward
    assert fake.multi == [2, 2]


class _TwoUses(torch.autograd.Function):
    """backward asks for gradient buffers the way the fused operators do"""
    seen = []

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(w)
        return x * 1.0

    @staticmethod
    def backward(ctx, dy):
        (w,) = ctx.saved_tensors
        (g,), acc = F._grads_like([w], N.MODE_BF16)
        _TwoUses.seen.append((type(g), g.data_ptr()))  # (no reference: a second owner would make autograd copy)
        out = F._out(dy, g)
        return out


def test_second_use_of_a_parameter_adds_into_the_first_buffer(fake):
    w = torch.nn.Parameter(torch.randn(5, 3))
    x = torch.randn(5, 3, requires_grad=True)
    _TwoUses.seen = []
    y = _TwoUses.apply(_TwoUses.apply(x, w), w)
    y.sum().backward()
    (t1, p1), (t2, p2) = _TwoUses.seen
    assert issubclass(t1, torch.Tensor) and t2 is F._Alias
    assert p2 == p1                                         # same memory, handed to autograd only once
    assert w.grad is not None and w.grad.data_ptr() == p1   # adopted without a copy or an add
    assert F._PASS.pending == {} and F._PASS.slabs == []                                 # forgotten when the backward pass ended


class _Frozen(torch.autograd.Function):
    """an operator whose parameter is frozen while its input still needs a gradient"""
    seen = []

    @staticmethod
    def forward(ctx, x, w):
        ctx.save_for_backward(w)
        return x * 1.0

    @staticmethod
    def backward(ctx, dy):
        (w,) = ctx.saved_tensors
        (g,), acc = F._grads_like([w], N.MODE_BF16, ctx.needs_input_grad[1:2])
        _Frozen.seen.append((type(g), g.data_ptr()))
        return F._out(dy, g)


def test_frozen_parameter_buffers_are_never_registered_or_returned(fake):
    """a frozen shared parameter's buffer dies with the backward that made it; a later operator must get
    fresh scratch, not an alias to freed memory, and nothing may come back to autograd for it."""
    w = torch.nn.Parameter(torch.randn(5, 3), requires_grad=False)
    x = torch.randn(5, 3, requires_grad=True)
    _Frozen.seen = []
    y = _Frozen.apply(_Frozen.apply(x, w), w)
    y.sum().backward()
    (t1, p1), (t2, p2) = _Frozen.seen
    assert issubclass(t1, torch.Tensor) and issubclass(t2, torch.Tensor)  # scratch both times, never an _Alias
    assert w.grad is None and x.grad is not None
    assert F._PASS.pending == {} and F._PASS.slabs == []


def test_direct_calls_outside_a_backward_pass_share_nothing(fake):
    w = torch.nn.Parameter(torch.randn(5, 3))
    (g1,), _ = F._grads_like([w], N.MODE_BF16)
    (g2,), _ = F._grads_like([w], N.MODE_BF16)
    assert isinstance(g1, torch.Tensor) and isinstance(g2, torch.Tensor) and g1.data_ptr() != g2.data_ptr()
    assert F._PASS.pending == {} and F._PASS.slabs == []


# ---------------------------------------------------------------------------------------------------------------
# Activation gradients with two consumers (functional._act_pending / _act_publish): the first backward publishes the
# buffer it returns, the second adds into it and returns None.  CPU stand-ins for the CUDA operators follow the
# protocol of KVProjectFn / CrossAttendFn / LatentBlockFn.backward.
# ---------------------------------------------------------------------------------------------------------------
class _Scale(torch.autograd.Function):
    """y = k x with the activation-gradient protocol of the fused operators"""

    @staticmethod
    def forward(ctx, x, k):
        ctx.save_for_backward(x)
        ctx.k = k
        return x * k

    @staticmethod
    @torch.autograd.function.once_differentiable
    def backward(ctx, dy):
        (x,) = ctx.saved_tensors
        merged = F._act_pending(x)
        if merged is not None:
            merged.add_(dy * ctx.k)  # (the kernels add inside the producing kernel: PIO_ACC_DX / PIO_ACC_DKV)
            return None, None
        dx = dy * ctx.k
        F._act_publish(x, dx)
        return dx, None


def test_second_consumer_of_an_activation_adds_into_the_first_gradient_buffer():
    F._ACT_STATS["merged"] = 0
    a = torch.randn(5, 3, requires_grad=True)
    x = a.exp()                                  # a non-leaf: identified by (grad_fn, output number, address, shape)
    y = _Scale.apply(x, 2.0).sum() + _Scale.apply(x, 3.0).sum()
    y.backward()
    assert F._ACT_STATS["merged"] == 1
    assert not F._PASS.acts                      # released by the end-of-pass callback
    torch.testing.assert_close(a.grad, a.detach().exp() * 5.0)
    # a second pass starts from an empty table
    a.grad = None
    x = a.exp()
    (_Scale.apply(x, 2.0).sum() + _Scale.apply(x, 3.0).sum() + _Scale.apply(x, 7.0).sum()).backward()
    assert F._ACT_STATS["merged"] == 3
    torch.testing.assert_close(a.grad, a.detach().exp() * 12.0)


def test_a_consumer_outside_the_protocol_never_costs_a_share_of_the_gradient():
    """autograd sums a foreign consumer's gradient with the published buffer into ANOTHER tensor; a share merged into
    the published buffer after that would be lost, so a producer merges only while the engine still holds the buffer:
    whichever order the engine runs the consumers in, the gradient is the full sum"""
    for variant in range(6):
        a = torch.randn(5, 3, requires_grad=True)
        x = a.exp()
        t = [_Scale.apply(x, 2.0).sum(), _Scale.apply(x, 3.0).sum()]
        f = (x * 5.0).sum() if variant < 3 else _Scale.apply(x.clone(), 5.0).sum()  # an ordinary torch op on the same activation
        y = (f + t[0] + t[1], t[0] + f + t[1], t[0] + t[1] + f)[variant % 3]
        y.backward()
        torch.testing.assert_close(a.grad, a.detach().exp() * 10.0)
        assert not F._PASS.acts


def test_leaves_and_direct_calls_are_never_merged(monkeypatch):
    F._ACT_STATS["merged"] = 0
    a = torch.randn(4, requires_grad=True)       # a leaf has no grad_fn to be identified by
    (_Scale.apply(a, 2.0).sum() + _Scale.apply(a, 3.0).sum()).backward()
    assert F._ACT_STATS["merged"] == 0
    torch.testing.assert_close(a.grad, torch.full((4,), 5.0))
    # outside a backward pass nothing can clear an entry, so nothing is published
    x = a.exp()
    F._act_publish(x, torch.zeros_like(x))
    assert not F._PASS.acts
    # the switch
    monkeypatch.setenv("PIO_NO_GRAD_MERGE", "1")
    a.grad = None
    x = a.exp()
    (_Scale.apply(x, 2.0).sum() + _Scale.apply(x, 3.0).sum()).backward()
    assert F._ACT_STATS["merged"] == 0
    torch.testing.assert_close(a.grad, a.detach().exp() * 5.0)


def test_same_memory_other_autograd_output_is_not_merged():
    """two views of one buffer are different autograd outputs: same address and shape, but not the same activation"""
    F._ACT_STATS["merged"] = 0
    a = torch.randn(6, requires_grad=True)
    x = a.exp()
    v1, v2 = x.view(6), x.view(6)                # same data_ptr, same shape, different grad_fn
    (_Scale.apply(v1, 2.0).sum() + _Scale.apply(v2, 3.0).sum()).backward()
    assert F._ACT_STATS["merged"] == 0
    torch.testing.assert_close(a.grad, a.detach().exp() * 5.0)
