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
