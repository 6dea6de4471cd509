"""Generates tests/golden/agg_dropout.pt: the UNMODIFIED reference aggregator (model/aggregator.py, imported from
/root/reference; never copied) in TRAINING mode with dropout > 0, together with the dropout masks torch drew.

torch's generator stream cannot be reproduced elsewhere, but WHERE dropout acts and how survivors are rescaled can be
pinned: `torch.nn.functional.dropout` and `torch.nn.functional.scaled_dot_product_attention` are wrapped (for the
duration of the run) by mathematically identical versions that draw their own Bernoulli masks with torch.rand and
RECORD them.  tests/test_oracle_golden.py replays the recorded masks through oracle/paths_oracle.py's hand-written
decoder (`RecordedDropout`) and must reproduce the module's output and gradients.  Run in the build container only:

    python tests/golden/make_golden_dropout.py
"""
import math
import os
import sys
import types

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PATHS_REFERENCE", "/root/reference")
for name in ["tiatoolbox", "tiatoolbox.wsicore", "tiatoolbox.wsicore.wsireader", "tiatoolbox.tools",
             "tiatoolbox.tools.tissuemask"]:
    sys.modules[name] = types.ModuleType(name)
sys.modules["tiatoolbox.wsicore.wsireader"].WSIReader = object
sys.modules["tiatoolbox.tools.tissuemask"].OtsuTissueMasker = object
sys.path.insert(0, REF)

import torch  # noqa: E402
import torch.nn.functional as F  # noqa: E402
from model.aggregator import TransformerAggregator  # noqa: E402

records = []          # masks of every dropout site that saw a non-empty tensor, in call order


def rec_dropout(input, p=0.5, training=True, inplace=False):
    if not training or p == 0.0:
        return input
    mask = (torch.rand_like(input) >= p).to(input.dtype) / (1.0 - p)
    if input.numel() > 0:
        records.append(("dropout", mask.clone()))
    return input * mask


def rec_sdpa(query, key, value, attn_mask=None, dropout_p=0.0, is_causal=False, scale=None, **kw):
    assert not is_causal
    s = query @ key.transpose(-1, -2) * (scale if scale is not None else 1.0 / math.sqrt(query.shape[-1]))
    if attn_mask is not None:
        s = s.masked_fill(attn_mask, float("-inf")) if attn_mask.dtype == torch.bool else s + attn_mask
    a = torch.softmax(s, dim=-1)
    if dropout_p > 0.0:
        mask = (torch.rand_like(a) >= dropout_p).to(a.dtype) / (1.0 - dropout_p)
        if a.numel() > 0:
            records.append(("attn", mask.clone()))
        a = a * mask
    return a @ value


def main():
    B, N, d, heads, layers, p = 3, 7, 32, 4, 2, 0.3
    torch.manual_seed(7)
    agg = TransformerAggregator(input_dim=48, model_dim=d, output_dim=d, nhead=heads, layers=layers, dropout=p)
    for q in agg.parameters():                                   # no zero-initialised biases: every term must show
        if q.dim() == 1:
            torch.nn.init.normal_(q, std=0.3)
    agg.train()
    tokens = torch.randn(B, N, d, requires_grad=True)
    num_ims = torch.tensor([7, 3, 5])
    enc_in = torch.zeros(B, 0, d)                                # model/paths.py:116

    orig = (F.dropout, F.scaled_dot_product_attention)
    F.dropout, F.scaled_dot_product_attention = rec_dropout, rec_sdpa
    try:
        out = agg(enc_in, tokens, None, num_ims)
    finally:
        F.dropout, F.scaled_dot_product_attention = orig
    w = torch.randn(B, d)
    (out * w).sum().backward()

    # per decoder layer, in call order: self-attention probabilities, dropout1, dropout2, FFN inner dropout, dropout3
    kinds = ["attn", "dropout", "dropout", "dropout", "dropout"]
    assert len(records) == 5 * layers and [k for k, _ in records] == kinds * layers, [k for k, _ in records]
    masks = {(j, site): records[5 * j + site][1] for j in range(layers) for site in range(5)}
    sd = {k: v.detach().clone() for k, v in agg.state_dict().items()}
    grads = {k: (v.grad.detach().clone() if v.grad is not None else None) for k, v in agg.named_parameters()}
    torch.save({"dims": dict(B=B, N=N, d=d, heads=heads, layers=layers, p=p), "state_dict": sd, "tokens": tokens.detach(),
                "num_ims": num_ims, "masks": masks, "out": out.detach(), "w": w, "grads": grads,
                "dtokens": tokens.grad.detach().clone()}, os.path.join(HERE, "agg_dropout.pt"))
    print("wrote agg_dropout.pt; out", out.detach().abs().mean().item(), "masks", {k: tuple(v.shape) for k, v in masks.items()})


if __name__ == "__main__":
    main()
