"""Container / helper semantics the model relies on (reference perceiver/utils.py)."""
import torch
import torch.nn as nn


class Sequential(nn.Sequential):
    """nn.Sequential whose stages may return tuples: a tuple is splatted into the next stage's
    positional arguments, anything else is passed as the single argument (reference utils.py:7-14).
    This is what lets ``PerceiverIO(x, pad_mask)`` hand ``(x_latent, pad_mask)`` to the decoder
    and why masks only ever reach the first (cross-attention) stage of a layer."""

    def forward(self, *inputs):
        for stage in self:
            inputs = stage(*inputs) if type(inputs) == tuple else stage(inputs)
        return inputs


def freeze(module: nn.Module):
    for p in module.parameters():
        p.requires_grad = False


def predict_masked_samples(masked_samples, encode_fn, tokenizer, model, num_predictions=5, device=None):
    """Top-k fill-in of [MASK] positions (reference utils.py:22-43); inference-time helper."""
    from .tokenizer import MASK_TOKEN
    xs, ms = encode_fn(masked_samples)
    xs, ms = xs.to(device), ms.to(device)
    with torch.no_grad():
        logits, _ = model(xs, ms, masking=False)
    at_mask = xs == tokenizer.token_to_id(MASK_TOKEN)
    _, top = torch.topk(logits[at_mask], k=num_predictions, dim=-1)
    filled = xs.clone()
    decoded = [[] for _ in masked_samples]
    for i in range(num_predictions):
        filled[at_mask] = top[:, i]
        for j in range(len(masked_samples)):
            decoded[j].append(tokenizer.decode(filled[j].tolist(), skip_special_tokens=True))
    return decoded
