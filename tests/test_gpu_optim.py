"""perceiver.FusedAdamW (pio_adamw_step) against torch.optim.AdamW: same update, bf16 shadows written with it."""
import pytest
import torch

pytestmark = pytest.mark.gpu
DEV = "cuda"
SHAPES = [(256, 256), (10003, 64), (3, 131), (517,), (1,), (40000,), (33, 7)]


def _params(seed):
    g = torch.Generator().manual_seed(seed)
    return [torch.nn.Parameter(torch.randn(*s, generator=g).to(DEV)) for s in SHAPES]


def test_matches_torch_adamw_over_several_steps_with_a_schedule():
    from perceiver import FusedAdamW
    pa, pb = _params(0), _params(0)
    kw = dict(lr=3e-3, betas=(0.9, 0.98), eps=1e-8, weight_decay=0.05)
    oa, ob = torch.optim.AdamW(pa, **kw), FusedAdamW(pb, **kw)
    sa = torch.optim.lr_scheduler.OneCycleLR(oa, max_lr=1e-2, total_steps=100)
    sb = torch.optim.lr_scheduler.OneCycleLR(ob, max_lr=1e-2, total_steps=100)
    g = torch.Generator().manual_seed(1)
    for it in range(7):
        for a, b in zip(pa, pb):
            grad = torch.randn(a.shape, generator=g).to(DEV) * (0.1 + it)
            a.grad, b.grad = grad.clone(), grad.clone()
        oa.step(); ob.step(); sa.step(); sb.step()
    for a, b, s in zip(pa, pb, SHAPES):
        torch.testing.assert_close(b, a, rtol=2e-5, atol=1e-6, msg=f"param {s}")
        torch.testing.assert_close(ob.state[b]["exp_avg"], oa.state[a]["exp_avg"], rtol=1e-5, atol=5e-6)  # |m| ~ 1: fp32 rounding of lerp vs b1*m + (1-b1)*g
        torch.testing.assert_close(ob.state[b]["exp_avg_sq"], oa.state[a]["exp_avg_sq"], rtol=1e-5, atol=5e-6)
    assert float(ob.state[pb[0]]["step"].item()) == 7.0


def test_skips_parameters_without_gradient_and_resumes_from_state_dict():
    from perceiver import FusedAdamW
    pa, pb = _params(2), _params(2)
    oa, ob = torch.optim.AdamW(pa, lr=1e-2), FusedAdamW(pb, lr=1e-2)
    g = torch.Generator().manual_seed(3)

    def grads(skip):
        for i, (a, b) in enumerate(zip(pa, pb)):
            if i == skip:
                a.grad = b.grad = None
                continue
            grad = torch.randn(a.shape, generator=g).to(DEV)
            a.grad, b.grad = grad.clone(), grad.clone()
    for it in range(3):
        grads(skip=1 if it == 1 else -1)
        if it == 1:
            continue  # (torch keeps per-tensor step counts, this optimizer one per group: only compare full steps)
        oa.step(); ob.step()
    import copy
    sd = copy.deepcopy(ob.state_dict())  # state_dict() hands out references to the live moment buffers
    pc = [torch.nn.Parameter(b.detach().clone()) for b in pb]
    oc = FusedAdamW(pc, lr=1e-2)
    oc.load_state_dict(sd)
    grads(-1)
    for b, c in zip(pb, pc):
        c.grad = b.grad.clone()
    oa.step(); ob.step(); oc.step()
    for a, b, c in zip(pa, pb, pc):
        torch.testing.assert_close(b, a, rtol=2e-5, atol=1e-6)
        torch.testing.assert_close(c, b, rtol=0, atol=0)


def test_writes_bf16_shadows_and_training_forward_skips_the_recast():
    import perceiver
    from perceiver import FusedAdamW, PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    from perceiver import _native as N
    perceiver.set_compute_mode("bf16")
    try:
        torch.manual_seed(0)
        enc = PerceiverEncoder(TextInputAdapter(503, 64, 72), (16, 64), 2, 4, 4, 2)  # 72 input channels: padded pitch stays 72
        dec = PerceiverDecoder(TextOutputAdapter(503, 64, num_output_channels=64), (16, 64), 4)
        model = PerceiverMLM(enc, dec, TextMasking(503, 1, 2, 3)).cuda()
        opt = FusedAdamW(model.parameters(), lr=1e-2)
        ids = torch.randint(3, 503, (8, 64), device=DEV)
        pad = torch.zeros(8, 64, dtype=torch.bool, device=DEV)
        losses = []
        for it in range(5):
            c0 = N.lib().pio_launch_count()
            loss, _ = model(ids.clone(), pad, return_loss=True)
            fwd_launches = N.lib().pio_launch_count() - c0
            shadowed = [p for p in model.parameters() if getattr(p, "_pio_w16", None) is not None]
            for p in shadowed:
                assert torch.equal(p._pio_w16[2][:, :p.shape[1]], p.detach().to(torch.bfloat16)), f"stale shadow, step {it}"
            if it == 2:
                ref_launches = fwd_launches
            if it > 2:
                assert fwd_launches == ref_launches  # steady state: no cast launch (it == 1 still had pio_cast_bf16_multi)
            opt.zero_grad(set_to_none=True)
            loss.backward()
            opt.step()
            losses.append(loss.item())
        assert losses[-1] < losses[0]
    finally:
        perceiver.set_compute_mode("fp32")


def test_step_inside_a_cuda_graph_advances_its_own_counter():
    from perceiver import FusedAdamW
    pa, pb = _params(5), _params(5)
    oa, ob = torch.optim.AdamW(pa, lr=2e-3), FusedAdamW(pb, lr=2e-3)
    static_g = [torch.zeros_like(b) for b in pb]
    for b, sg in zip(pb, static_g):
        b.grad = sg
    g = torch.Generator().manual_seed(6)

    def new_grads():
        for a, sg in zip(pa, static_g):
            grad = torch.randn(a.shape, generator=g).to(DEV)
            a.grad = grad.clone()
            sg.copy_(grad)
    side = torch.cuda.Stream()
    side.wait_stream(torch.cuda.current_stream())
    with torch.cuda.stream(side):
        new_grads(); oa.step(); ob.step()
    torch.cuda.current_stream().wait_stream(side)
    graph = torch.cuda.CUDAGraph()
    new_grads(); oa.step()
    with torch.cuda.graph(graph):
        ob.step()
    graph.replay()
    for _ in range(3):
        new_grads(); oa.step(); graph.replay()
    torch.cuda.synchronize()
    for a, b in zip(pa, pb):
        torch.testing.assert_close(b, a, rtol=2e-5, atol=1e-6)
    assert float(ob.state[pb[0]]["step"].item()) == 5.0


def test_flat_gather_and_scatter_around_the_gradient_all_reduce():
    """pio_flat_copy_f32 through perceiver.distributed._native_plan: gather == torch.cat, scatter(scale) == views * scale."""
    from perceiver import _native as N
    from perceiver import distributed as D
    params = _params(9)
    g = torch.Generator().manual_seed(10)
    for p in params:
        p.grad = torch.randn(p.shape, generator=g).to(DEV)
    plan = D._native_plan(params)
    assert plan is not None
    L, st = plan["lib"], plan["stream"]
    N.check(L.pio_flat_copy_f32(plan["flat"].data_ptr(), plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], 1, 1.0, st), "gather")
    torch.cuda.synchronize()
    for p, off in zip(params, plan["offs"]):
        assert torch.equal(plan["flat"][off:off + p.numel()], p.grad.reshape(-1))
    want = [p.grad.clone() * 0.25 for p in params]
    for p in params:
        p.grad.fill_(float("nan"))
    N.check(L.pio_flat_copy_f32(plan["flat"].data_ptr(), plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], 0, 0.25, st), "scatter")
    torch.cuda.synchronize()
    for p, w in zip(params, want):
        assert torch.equal(p.grad, w)
    # new gradient buffers (eager training): the pointer table follows
    for p in params:
        p.grad = torch.ones_like(p)
    plan2 = D._native_plan(params)
    assert plan2 is plan
    N.check(L.pio_flat_copy_f32(plan["flat"].data_ptr(), plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], 1, 1.0, st), "gather")
    torch.cuda.synchronize()
    assert all(bool((plan["flat"][off:off + p.numel()] == 1).all()) for p, off in zip(params, plan["offs"]))
