"""AdamW as one kernel launch per step (pio_adamw_step), drop-in for ``torch.optim.AdamW``.

The reference's LightningModules step ``torch.optim.AdamW`` (lightning.py:67-72).  This optimizer applies the
same update (checked against torch in tests/test_gpu_optim.py) with one multi-tensor launch that also writes
the bf16 weight shadows the tensor-core GEMMs read, so in bf16 mode the next forward does not re-cast them.
Learning-rate schedulers work as usual (they edit ``param_groups``; the values are mirrored into a small
device buffer the kernel reads), and a step can be captured in a CUDA graph: the step counter lives on the device.
"""
import numpy as np
import torch

from . import _native as N

_CHUNK = 16384  # PIO_ADAMW_CHUNK


class FusedAdamW(torch.optim.Optimizer):
    def __init__(self, params, lr=1e-3, betas=(0.9, 0.999), eps=1e-8, weight_decay=1e-2):
        if lr < 0.0 or eps < 0.0 or weight_decay < 0.0 or not (0.0 <= betas[0] < 1.0 and 0.0 <= betas[1] < 1.0):
            raise ValueError("invalid AdamW hyper-parameters")
        super().__init__(params, dict(lr=lr, betas=betas, eps=eps, weight_decay=weight_decay))
        self._plans = {}

    # ------------------------------------------------------------------ per-group launch plan
    def _plan(self, gi, group, params):
        """chunk table (built once per parameter list), tensor table (rewritten when a pointer moved), hyper buffer"""
        plan = self._plans.get(gi)
        dev = params[0].device
        ident = tuple(id(p) for p in params)
        if plan is None or plan["ident"] != ident:
            for p in params:
                N.require_cuda(p)
                if p.dtype != torch.float32 or not p.is_contiguous():
                    raise TypeError("FusedAdamW: parameters must be contiguous fp32 CUDA tensors")
                st = self.state[p]
                if "exp_avg" not in st:
                    st["exp_avg"] = torch.zeros_like(p)
                    st["exp_avg_sq"] = torch.zeros_like(p)
            chunks = []
            for ti, p in enumerate(params):
                n = p.numel()
                for off in range(0, n, _CHUNK):
                    chunks.append((ti | (min(_CHUNK, n - off) << 32), off))  # {int32 tensor, int32 count, int64 offset}
            chunk_t = torch.from_numpy(np.asarray(chunks, dtype=np.int64).reshape(-1, 2)).to(dev)
            steps = 0.0
            if plan is not None:
                steps = float(plan["hyper"][5].item())
            else:
                prev = [self.state[p].get("step") for p in params if "step" in self.state[p]]
                if prev:  # resumed from a state_dict
                    steps = float(prev[0]) if not torch.is_tensor(prev[0]) else float(prev[0].item())
            hyper = torch.zeros(6, dtype=torch.float32, device=dev)
            hyper[5] = steps
            plan = dict(ident=ident, chunks=chunk_t, n_chunks=len(chunks), hyper=hyper, hyper_host=None, key=None,
                        table=torch.empty(len(params), 6, dtype=torch.int64, device=dev))
            for p in params:
                self.state[p]["step"] = hyper[5:6]  # one device counter per group, shared by its tensors
            self._plans[gi] = plan
        # tensor table
        rows = []
        for p in params:
            st = self.state[p]
            sh = getattr(p, "_pio_w16", None)
            shadow = sh[2] if (sh is not None and p.dim() == 2 and sh[1] == p.data_ptr() and sh[2].device == dev) else None
            rows.append((p.data_ptr(), p.grad.data_ptr(), st["exp_avg"].data_ptr(), st["exp_avg_sq"].data_ptr(),
                         shadow.data_ptr() if shadow is not None else 0,
                         (p.shape[1] | (shadow.shape[1] << 32)) if shadow is not None else 0))
        key = tuple(rows)
        if key != plan["key"]:
            for p in params:
                g = p.grad
                if g.dtype != torch.float32 or not g.is_contiguous() or g.device != dev:
                    raise TypeError("FusedAdamW: gradients must be contiguous fp32 tensors on the parameters' device")
            if torch.cuda.is_current_stream_capturing():
                raise RuntimeError("FusedAdamW: parameter / gradient / shadow addresses changed while the stream is being "
                                   "captured; run one eager step with the final buffers first")
            # a plain synchronous copy from pageable memory: the host array may be rewritten as soon as this returns
            # (a pinned staging buffer + non_blocking copy could be overwritten by the next step's table while the
            # previous copy was still in flight — eager training hands out new gradient buffers every step)
            plan["table"].copy_(torch.from_numpy(np.asarray(rows, dtype=np.int64)))
            plan["key"] = key
            plan["shadowed"] = [p for p, r in zip(params, rows) if r[4] != 0]
        # hyper-parameters (the step counter stays on the device)
        hp = (float(group["lr"]), float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]),
              float(group["weight_decay"]))
        if hp != plan["hyper_host"]:
            plan["hyper"][:5].copy_(torch.tensor(hp, dtype=torch.float32))
            plan["hyper_host"] = hp
        return plan

    def sync_hyper(self):
        """Mirror ``param_groups`` (lr, betas, eps, weight_decay — e.g. after ``scheduler.step()``) into the device
        buffers the kernel reads.  ``step()`` does this itself; a step REPLAYED from a CUDA graph does not run any
        Python, so call this before ``graph.replay()`` whenever a scheduler changed the values."""
        for key, plan in self._plans.items():
            gi = key[1] if isinstance(key, tuple) else key  # (subset key, group index) for step_params plans
            group = self.param_groups[gi]
            hp = (float(group["lr"]), float(group["betas"][0]), float(group["betas"][1]), float(group["eps"]),
                  float(group["weight_decay"]))
            if hp != plan["hyper_host"]:
                plan["hyper"][:5].copy_(torch.tensor(hp, dtype=torch.float32))
                plan["hyper_host"] = hp

    def load_state_dict(self, state_dict):
        """resume: the launch plans (and their device step counters) are rebuilt from the loaded state"""
        super().load_state_dict(state_dict)
        self._plans = {}

    def _launch(self, key, group, params):
        L = N.lib()
        plan = self._plan(key, group, params)
        st = N.stream_ptr(params[0].device)
        N.check(L.pio_adamw_step(plan["table"].data_ptr(), plan["chunks"].data_ptr(), plan["n_chunks"], plan["hyper"].data_ptr(), st),
                "pio_adamw_step")
        N.check(L.pio_adamw_tick(plan["hyper"].data_ptr(), st), "pio_adamw_tick")
        for p in plan["shadowed"]:  # their bf16 shadows are current: the next training forward need not re-cast them
            p._pio_fresh16 = True

    @torch.no_grad()
    def step(self, closure=None):
        loss = None
        if closure is not None:
            with torch.enable_grad():
                loss = closure()
        for gi, group in enumerate(self.param_groups):
            params = [p for p in group["params"] if p.grad is not None]
            if params:
                self._launch(gi, group, params)
        return loss

    @torch.no_grad()
    def step_params(self, params, key):
        """The same update for a SUBSET of the parameters, on the current stream: one launch per parameter group that
        has members in ``params``.  ``key`` names the subset (its launch plan — chunk table, pointer table, device
        step counter — is kept under it), so call it with the same subsets every step: `GraphedTrainStep` steps each
        gradient bucket of `FlatGradients` as soon as the backward pass has completed (and all-reduced) it, instead of
        one launch over everything after the pass.  Do not mix with ``step()`` on the same parameters within one run:
        every plan counts its own steps."""
        wanted = {id(p) for p in params}
        for gi, group in enumerate(self.param_groups):
            sub = [p for p in group["params"] if id(p) in wanted and p.grad is not None]
            if sub:
                self._launch((key, gi), group, sub)
