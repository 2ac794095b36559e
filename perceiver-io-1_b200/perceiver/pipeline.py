"""Input pipelining for the MLM train step.

TextMasking (reference model.py:265-293) needs one host read per batch — the length of its
``torch.randint`` draw is the number of random-replacement positions — so a step that masks its own batch
leaves the GPU idle while the host waits for that number and then launches the step.  `MaskingPrefetcher`
masks batch i+1 on a side stream while step i computes: same kernels, same torch RNG draws in the same
order (so the same generator state gives bit-identical ids / labels), only earlier.
"""
import torch


class MaskingPrefetcher:
    """submit(ids, pad) -> copies the batch (host or device) into staging buffers and masks it on a side
    stream; take(out_ids, out_labels[, out_pad]) -> makes the caller's stream wait for that work and copies
    the result into the caller's (e.g. CUDA-graph static) buffers.  One batch in flight."""

    def __init__(self, masking, device):
        self.masking = masking
        self.device = torch.device(device)
        self.side = torch.cuda.Stream(self.device)
        self.ready = torch.cuda.Event()
        self.free = torch.cuda.Event()
        self.ids = self.labels = self.pad = None
        self.pending = False

    def submit(self, ids, pad_mask, generator=None, after_current_stream=False):
        """``after_current_stream=True`` orders the side stream behind everything already queued on the
        caller's stream (needed only when ``ids`` / ``pad_mask`` were just produced there; it also makes the
        masking wait for a train step queued before it, which defeats the overlap)."""
        if self.pending:
            raise RuntimeError("MaskingPrefetcher: take() the submitted batch before submitting the next one")
        if self.ids is None or self.ids.shape != ids.shape:
            self.ids = torch.empty(ids.shape, dtype=torch.int64, device=self.device)
            self.labels = torch.empty_like(self.ids)
            self.pad = torch.empty(pad_mask.shape, dtype=torch.bool, device=self.device) if pad_mask is not None else None
        if after_current_stream:
            self.side.wait_stream(torch.cuda.current_stream(self.device))
        with torch.cuda.stream(self.side):
            self.side.wait_event(self.free)  # the previous batch has been copied out of the staging buffers
            self.ids.copy_(ids, non_blocking=True)
            if pad_mask is not None:
                self.pad.copy_(pad_mask, non_blocking=True)
            _, labels = self.masking(self.ids, self.pad if pad_mask is not None else None, generator=generator)
            self.labels.copy_(labels, non_blocking=True)
            self.ready.record(self.side)
        self.pending = True

    def take(self, out_ids, out_labels, out_pad=None):
        if not self.pending:
            raise RuntimeError("MaskingPrefetcher: nothing submitted")
        cur = torch.cuda.current_stream(self.device)
        cur.wait_event(self.ready)
        out_ids.copy_(self.ids, non_blocking=True)
        out_labels.copy_(self.labels, non_blocking=True)
        if out_pad is not None and self.pad is not None:
            out_pad.copy_(self.pad, non_blocking=True)
        self.free.record(cur)
        self.pending = False


class GraphedTrainStep:
    """forward + loss + backward + optimizer step of any model built from this package's operators as ONE replayed CUDA
    graph.  ``loss_fn()`` computes the loss from tensors at fixed addresses (copy each batch into them before calling
    ``replay()``); parameter gradients are written in place into one flat buffer (`FlatGradients`), and with an
    initialised process group of more than one rank its buckets are all-reduced (mean) on a side stream from inside the
    captured backward pass.  ``overlap_allreduce=False``: one flat all-reduce and the optimizer run eagerly after the
    graph (the previous scheme).  bf16 compute mode; the optimizer must be capturable (`perceiver.FusedAdamW`).
    Dropout seeds are launch arguments, so a replay repeats the captured masks: fine for timing, not for training with
    dropout > 0 (train those eagerly).
    With a process group of more than one rank, construction and every replay are COLLECTIVE: all ranks must build the
    step and replay it the same number of times (a step built on one rank alone waits for the others in its first
    all-reduce forever — bench.py's side measurements did exactly that once)."""

    def __init__(self, params, optimizer, loss_fn, overlap_allreduce=True, bucket_bytes=16 << 20, warmup=3, before_warmup_step=None,
                 bucketed_optimizer=None):
        import torch.distributed as dist
        from . import _native as N
        from .distributed import FlatGradients, average_gradients
        self.params = [p for p in params if p.requires_grad]
        dev = self.params[0].device
        self.opt, self.device, self.loss_fn = optimizer, dev, loss_fn
        self.world = dist.get_world_size() if (dist.is_available() and dist.is_initialized()) else 1
        use_fg = overlap_allreduce or self.world == 1
        self.fg = FlatGradients(self.params, bucket_bytes=bucket_bytes) if use_fg else None
        self._average = average_gradients
        # ``bucketed_optimizer`` (off unless asked for, or PIO_BUCKET_ADAMW=1): the optimizer steps each gradient bucket on its
        # own lowest-priority stream as soon as the backward pass has completed it — and, with more than one rank, its
        # all-reduce has finished — instead of one launch over all parameters after the pass, so that only the last
        # bucket's update is left at the end of the step.  Measured at one rank on cfg2: 1.833 ms against 1.824 ms with the
        # single launch — some stream keeps every SM busy for 98 % of the step, and AdamW's 30 bytes per parameter of HBM
        # traffic slow the kernels they overlap by as much as they save at the end; kept for many-rank runs, where the
        # update otherwise waits for the last all-reduce.
        import os as _os
        if bucketed_optimizer is None:
            bucketed_optimizer = _os.environ.get("PIO_BUCKET_ADAMW") is not None
        nccl_or_single = self.world == 1 or dist.get_backend() == "nccl"
        self.bucketed_opt = bool(bucketed_optimizer) and self.fg is not None and hasattr(optimizer, "step_params") and nccl_or_single
        self.opt_stream = torch.cuda.Stream(dev) if self.bucketed_opt else None
        if self.bucketed_opt:
            self.fg.on_ready = self._bucket_ready
        side = torch.cuda.Stream(dev)
        side.wait_stream(torch.cuda.current_stream(dev))
        with torch.cuda.stream(side):
            if self.fg is not None:  # one ordinary pass teaches FlatGradients the completion order of the parameters
                if before_warmup_step is not None:
                    before_warmup_step()
                with self.fg.profile():
                    loss_fn().backward()
            for _ in range(warmup):
                if before_warmup_step is not None:
                    before_warmup_step()
                if self.fg is None:
                    optimizer.zero_grad(set_to_none=True)
                self._body()
                self._tail()
        torch.cuda.current_stream(dev).wait_stream(side)
        if before_warmup_step is not None:
            before_warmup_step()
        if self.fg is None:
            optimizer.zero_grad(set_to_none=True)
        self.graph = torch.cuda.CUDAGraph()
        before = N.lib().pio_launch_count()
        # The step is captured on a HIGH-priority stream: the library's side streams (parameter-gradient kernels) have the
        # default, lowest priority, so when SMs free up the pending blocks of the main chain are placed first — after a
        # latent block's backward chain the batched weight-gradient and LayerNorm-gradient kernels would otherwise keep
        # the next operator's kernels waiting.
        import os
        prio = None if os.environ.get("PIO_NO_CAPTURE_PRIORITY") else -1
        cap_stream = torch.cuda.Stream(dev, priority=prio) if prio is not None else torch.cuda.Stream(dev)
        with torch.cuda.graph(self.graph, stream=cap_stream, capture_error_mode="thread_local" if self.world > 1 else "global"):
            self.loss = self._body()
        self.launches_per_replay = N.lib().pio_launch_count() - before  # kernels of this library inside one replay

    @property
    def in_graph_optimizer(self):
        return self.fg is not None or self.world == 1

    def _bucket_ready(self, b, handle):
        if self.fg.flat is None or getattr(self, "_in_body", False) is False:
            return  # (the profiling pass, or a backward pass outside the step)
        cur = torch.cuda.current_stream(self.device)
        ev = torch.cuda.Event()
        ev.record(cur)  # the bucket's gradients are final — and nothing enqueued so far still reads its parameters' old values
        self.opt_stream.wait_event(ev)
        with torch.cuda.stream(self.opt_stream):
            if handle is not None:
                handle.wait()  # this stream waits for the bucket's all-reduce
            self.opt.step_params(self.fg.bucket_params(b), ("bucket", b))

    def _body(self):
        if self.fg is not None:
            self.fg.step_begin()
        self._in_body = True
        try:
            loss = self.loss_fn()
            loss.backward()
            if self.fg is not None:
                self.fg.step_end()
        finally:
            self._in_body = False
        if self.bucketed_opt:
            torch.cuda.current_stream(self.device).wait_stream(self.opt_stream)
        elif self.in_graph_optimizer:
            self.opt.step()
        return loss

    def _tail(self):
        if not self.in_graph_optimizer:  # previous scheme: exchange + optimizer eagerly after the graph
            self._average(self.params)
            self.opt.step()

    def replay(self):
        self.graph.replay()
        self._tail()
        return self.loss


class GraphedMLMStep(GraphedTrainStep):
    """The whole masked-LM train step of LitMaskedLanguageModel (reference lightning.py:219-231: masking, forward,
    cross-entropy, backward; then the optimizer step) as ONE replayed CUDA graph per batch.

        step = GraphedMLMStep(model, optimizer, ids, pad)     # model: PerceiverMLM on a CUDA device, bf16 mode
        for ids, pad in batches:                               # host (pinned) or device tensors of the same shape
            loss = step(ids, pad)                              # a device scalar (valid until the next call)

    * TextMasking runs outside the graph on a side stream, one batch ahead (`MaskingPrefetcher`): its one host read
      (the length of the reference's randint draw) no longer idles the GPU, results are bit-identical.
    * Inside the graph (`GraphedTrainStep`): forward with the fused vocabulary-projection + cross-entropy head,
      backward with every parameter gradient written in place into one flat buffer, the optimizer step, and for
      more than one rank the bucketed gradient all-reduce overlapped with the backward pass."""

    def __init__(self, model, optimizer, ids, pad, overlap_allreduce=True, bucket_bytes=16 << 20, warmup=3, bucketed_optimizer=None):
        dev = next(model.parameters()).device
        self.model = model
        self.ids = torch.empty(ids.shape, dtype=torch.int64, device=dev)
        self.labels = torch.empty_like(self.ids)
        self.pad = torch.empty(pad.shape, dtype=torch.bool, device=dev)
        self.prefetch = MaskingPrefetcher(model.masking, dev)

        def loss_fn():
            return model(self.ids, self.pad, masking=False, return_loss=True, labels=self.labels)[0]
        super().__init__(model.parameters(), optimizer, loss_fn, overlap_allreduce=overlap_allreduce, bucket_bytes=bucket_bytes,
                         warmup=warmup, before_warmup_step=lambda: self._fill(ids, pad), bucketed_optimizer=bucketed_optimizer)

    def _fill(self, ids, pad):
        if not self.prefetch.pending:
            self.prefetch.submit(ids, pad)
        self.prefetch.take(self.ids, self.labels, self.pad)

    def __call__(self, ids, pad):
        self._fill(ids, pad)        # consume the prefetched batch (the first call masks its own)
        loss = self.replay()
        self.prefetch.submit(ids, pad)  # next call's batch: H2D (for host tensors) + masking, overlapped with this step
        return loss
