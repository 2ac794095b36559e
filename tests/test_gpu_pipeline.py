"""MaskingPrefetcher: masking a batch ahead of time on a side stream gives bit-identical ids / labels."""
import pytest
import torch

pytestmark = pytest.mark.gpu


def _masking():
    from perceiver import TextMasking
    return TextMasking(vocab_size=10003, unk_token_id=1, mask_token_id=2, num_special_tokens=3)


@pytest.mark.parametrize("host_input", [False, True])
def test_prefetched_masking_is_bit_identical(host_input):
    from perceiver import MaskingPrefetcher
    dev = torch.device("cuda")
    g = torch.Generator().manual_seed(5)
    ids = torch.randint(3, 10003, (16, 512), generator=g)
    pad = torch.rand(16, 512, generator=g) > 0.9
    masking = _masking()
    # direct
    torch.manual_seed(123)
    ids_a = ids.to(dev)
    _, labels_a = masking(ids_a, pad.to(dev))
    # prefetched, two batches back to back through the same staging buffers
    torch.manual_seed(123)
    pf = MaskingPrefetcher(masking, dev)
    src_ids, src_pad = (ids.pin_memory(), pad.pin_memory()) if host_input else (ids.to(dev), pad.to(dev))
    torch.cuda.synchronize()
    out_ids, out_labels, out_pad = torch.empty_like(ids_a), torch.empty_like(ids_a), torch.empty(16, 512, dtype=torch.bool, device=dev)
    pf.submit(src_ids, src_pad)
    with pytest.raises(RuntimeError):
        pf.submit(src_ids, src_pad)
    pf.take(out_ids, out_labels, out_pad)
    torch.cuda.synchronize()
    assert torch.equal(out_ids, ids_a) and torch.equal(out_labels, labels_a) and torch.equal(out_pad.cpu(), pad)
    assert (out_labels != -100).any() and (out_ids != ids.to(dev)).any()
    # second batch: continues the RNG stream exactly like a second direct call would
    ids_b = ids.to(dev)
    pf.submit(src_ids, src_pad)
    pf.take(out_ids, out_labels)
    torch.cuda.synchronize()
    torch.manual_seed(123)
    _ = masking(ids.to(dev), pad.to(dev))
    _, labels_b = masking(ids_b, pad.to(dev))
    assert torch.equal(out_ids, ids_b) and torch.equal(out_labels, labels_b)
    with pytest.raises(RuntimeError):
        pf.take(out_ids, out_labels)
