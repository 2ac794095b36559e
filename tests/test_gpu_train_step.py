"""The train step exactly as bench.py runs it (perceiver.GraphedMLMStep: masking prefetched on a side stream,
CUDA-graph replay of forward + fused vocabulary/CE head + backward into flat gradient buffers + FusedAdamW) against
the eager step of the same model and, at step 1, against the fp32 oracle (reference lightning.py:219-231)."""
import copy
import os
import subprocess
import sys

import pytest
import torch

from gpu_util import BF16_TOL, assert_rel, bf16_grad_bound, grad_table, rel_err
from oracle import perceiver_oracle as O

pytestmark = pytest.mark.gpu
DEV = "cuda"
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


@pytest.fixture(autouse=True)
def bf16_mode():
    import perceiver
    perceiver.set_compute_mode("bf16")
    yield
    perceiver.set_compute_mode("fp32")
    from perceiver import functional as F
    F._DST.clear()


def _mlm(vocab, seq, latent, layers, sa, heads):
    from perceiver import PerceiverDecoder, PerceiverEncoder, PerceiverMLM, TextInputAdapter, TextMasking, TextOutputAdapter
    c = latent[1]
    enc = PerceiverEncoder(TextInputAdapter(vocab, seq, c), latent, layers, heads[0], heads[1], sa)
    dec = PerceiverDecoder(TextOutputAdapter(vocab, seq, c), latent, heads[2])
    return PerceiverMLM(enc, dec, TextMasking(vocab, 1, 2, 3))


def _batch(seed, batch, seq, vocab):
    g = torch.Generator().manual_seed(seed)
    ids = torch.randint(3, vocab, (batch, seq), generator=g)
    lens = torch.randint(seq // 4, seq + 1, (batch,), generator=g)
    lens[0] = seq
    pad = torch.arange(seq)[None, :] >= lens[:, None]
    ids[pad] = 0
    return ids, pad


def _global_rel(got, ref):
    num = sum(((got[k].double().cpu() - ref[k].double().cpu()) ** 2).sum().item() for k in ref)
    den = sum((ref[k].double() ** 2).sum().item() for k in ref)
    return (num / max(den, 1e-300)) ** 0.5


def test_flat_gradients_equal_the_per_operator_slabs_and_save_launches():
    import perceiver
    from perceiver import _native as N
    torch.manual_seed(0)
    model = _mlm(1003, 128, (32, 64), 3, 2, (2, 2, 2)).to(DEV)
    ids, pad = _batch(1, 8, 128, 1003)
    ids, pad = ids.to(DEV), pad.to(DEV)
    labels = torch.full_like(ids, -100)
    sel = (torch.rand(ids.shape, device=DEV) < 0.15) & ~pad
    labels[sel] = ids[sel]

    def run():
        c0 = N.lib().pio_launch_count()
        loss, _ = model(ids.clone(), pad, masking=False, return_loss=True, labels=labels)
        loss.backward()
        torch.cuda.synchronize()
        return loss.detach().clone(), N.lib().pio_launch_count() - c0

    model.zero_grad(set_to_none=True)
    loss_ref, n_ref = run()
    ref = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    fg = perceiver.FlatGradients(model.parameters(), bucket_bytes=64 << 10)
    with fg.profile():
        run()
    assert len(fg.views) >= 2
    for _ in range(2):
        fg.step_begin()
        loss, n_flat = run()
        fg.step_end()
    torch.cuda.synchronize()
    lo, hi = fg.flat.data_ptr(), fg.flat.data_ptr() + fg.flat.numel() * 4
    for k, p in model.named_parameters():
        assert lo <= p.grad.data_ptr() < hi, k
    got = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
    assert_rel(loss, loss_ref, 1e-6, "loss")
    assert _global_rel(got, ref) < 1e-5  # same kernels, same atomics: only the summation order may differ
    for k in ref:
        assert rel_err(got[k], ref[k]) < 1e-3, k
    assert all((fg.fired >> b) & 1 for b in range(fg.n_touched_buckets))  # every bucket was reported complete by the operators
    # an optimizer step on the views
    opt = perceiver.FusedAdamW(model.parameters(), lr=1e-3)
    before = {k: p.detach().clone() for k, p in model.named_parameters()}
    opt.step()
    torch.cuda.synchronize()
    assert any(not torch.equal(before[k], p.detach()) for k, p in model.named_parameters())
    fg.close()


@pytest.mark.parametrize("cfg", ["small", "cfg2_b64", "small_bucketed_adamw"])
def test_graphed_step_matches_eager_steps_and_the_oracle(cfg):
    """3 optimizer steps: CUDA-graph replay (prefetched masking, flat gradients, AdamW in the graph) against the eager
    step on a copy of the model fed the same masked batches; step 1 loss and gradients against the fp32 oracle."""
    import perceiver
    bucketed = cfg.endswith("_bucketed_adamw")  # AdamW per gradient bucket from inside the backward pass (two buckets here)
    if cfg.startswith("small"):
        vocab, seq, latent, layers, sa, heads, batch = 1003, 128, (32, 64), 3, 2, (2, 2, 2), 8
    else:  # BASELINE.json configs[1], the benchmarked workload
        vocab, seq, latent, layers, sa, heads, batch = 10003, 512, (64, 256), 3, 6, (4, 4, 4), 64
    torch.manual_seed(0)
    model = _mlm(vocab, seq, latent, layers, sa, heads).to(DEV)
    eager = copy.deepcopy(model)
    ids, pad = _batch(3, batch, seq, vocab)
    ids_d, pad_d = ids.to(DEV), pad.to(DEV)
    lr = 1e-3
    opt_g = perceiver.FusedAdamW(model.parameters(), lr=lr)
    opt_e = perceiver.FusedAdamW(eager.parameters(), lr=lr)
    torch.manual_seed(11)
    init = {k: p.detach().clone() for k, p in model.named_parameters()}
    gstep = perceiver.GraphedMLMStep(model, opt_g, ids_d, pad_d, bucketed_optimizer=bucketed,
                                     bucket_bytes=(256 << 10) if bucketed else (16 << 20))
    assert gstep.bucketed_opt == bucketed and (not bucketed or len(gstep.fg.views) >= 2)
    # construction ran warm-up optimizer steps: back to the initial weights and optimizer state
    with torch.no_grad():
        for k, p in model.named_parameters():
            p.copy_(init[k])
        for st in opt_g.state.values():
            st["exp_avg"].zero_()
            st["exp_avg_sq"].zero_()
        for plan in opt_g._plans.values():
            plan["hyper"][5] = 0.0
    assert perceiver.functional.refresh_weight_shadows(model)  # in place: the graph holds the shadow addresses
    losses_g, losses_e = [], []
    for it in range(3):
        loss = gstep(ids_d, pad_d)
        torch.cuda.synchronize()
        losses_g.append(loss.item())
        masked, labels = gstep.ids.clone(), gstep.labels.clone()  # the batch this replay consumed
        if it == 0:
            grads0 = {k: p.grad.detach().clone() for k, p in model.named_parameters()}
        opt_e.zero_grad(set_to_none=True)
        loss_e, _ = eager(masked, pad_d, masking=False, return_loss=True, labels=labels)
        loss_e.backward()
        if it == 0:
            grads0_e = {k: p.grad.detach().clone() for k, p in eager.named_parameters()}
        opt_e.step()
        losses_e.append(loss_e.item())
    torch.cuda.synchronize()
    for a, b in zip(losses_g, losses_e):
        assert abs(a - b) <= 2e-3 * abs(b), (losses_g, losses_e)
    assert _global_rel(grads0, grads0_e) < 2e-3
    final_g = {k: p.detach() for k, p in model.named_parameters()}
    final_e = {k: p.detach() for k, p in eager.named_parameters()}
    moved = _global_rel(final_e, init)
    assert moved > 1e-4
    # same trajectory: the two runs differ only in the order of the atomic gradient sums, but AdamW normalises every element's
    # update by its own gradient magnitude, so elements with a near-zero gradient can step differently: a tenth of the
    # distance travelled bounds that; a wrong gradient or a stale weight shadow moves the runs apart by the full distance
    assert _global_rel(final_g, final_e) < 0.1 * moved, (_global_rel(final_g, final_e), moved)


def _oracle_step(sd, masked, pad, labels, layers, sa, heads):
    logits = O.mlm_forward(sd, masked, pad, num_layers=layers, cross_heads=heads[0], self_heads=heads[1], sa_layers=sa,
                           dec_heads=heads[2])
    loss = O.cross_entropy(logits, labels)
    loss.backward()
    return loss


@pytest.mark.parametrize("cfg", ["small", "cfg2_b64"])
def test_graphed_step_first_step_matches_the_oracle(cfg):
    import perceiver
    if cfg == "small":
        vocab, seq, latent, layers, sa, heads, batch = 1003, 128, (32, 64), 3, 2, (2, 2, 2), 8
    else:
        vocab, seq, latent, layers, sa, heads, batch = 10003, 512, (64, 256), 3, 6, (4, 4, 4), 64
    torch.manual_seed(0)
    model = _mlm(vocab, seq, latent, layers, sa, heads).to(DEV)
    sd = {k: v.detach().cpu().clone().requires_grad_(v.is_floating_point()) for k, v in model.state_dict().items()}
    ids, pad = _batch(5, batch, seq, vocab)
    ids_d, pad_d = ids.to(DEV), pad.to(DEV)
    opt = perceiver.FusedAdamW(model.parameters(), lr=0.0, weight_decay=0.0)  # lr 0: warm-up steps leave the weights alone
    gstep = perceiver.GraphedMLMStep(model, opt, ids_d, pad_d)
    loss = gstep(ids_d, pad_d)
    torch.cuda.synchronize()
    masked, labels = gstep.ids.cpu(), gstep.labels.cpu()
    ref_loss = _oracle_step(sd, masked, pad, labels, layers, sa, heads)
    assert_rel(loss, ref_loss, BF16_TOL, "loss")
    got = {k: p.grad.detach() for k, p in model.named_parameters()}
    ref = {k: sd[k].grad for k in got if sd[k].grad is not None}
    rows, glob = grad_table(got, ref)
    print(f"whole-gradient rel L2 {glob:.3e}; worst tensors (max-rel, L2-rel, numel):")
    for k, e, l2, n in rows[:12]:
        print(f"   {e:.3e} {l2:.3e} {n:8d} {k}")
    assert glob < BF16_TOL
    bad = [(k, e) for k, e, _, _ in rows if e > bf16_grad_bound(k)]
    assert not bad, bad


@pytest.mark.timeout(600)
def test_two_gpu_overlapped_allreduce_equals_single_process_full_batch():
    """torchrun, 2 ranks: rank-averaged gradients from the bucketed, overlapped all-reduce inside the captured step ==
    single-process gradients on the concatenated batch."""
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    import socket
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    r = subprocess.run([sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node", "2", "--master-addr", "127.0.0.1",
                        "--master-port", str(port), os.path.join(ROOT, "tests", "ddp_worker.py")],
                       capture_output=True, text=True, timeout=540)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "DDP_WORKER_OK" in r.stdout, r.stdout[-3000:]
