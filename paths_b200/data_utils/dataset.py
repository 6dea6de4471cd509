"""collate_fn with the reference's output schema (data_utils/dataset.py:206-243).  Dataset / split loading
(SlideDataset, load_splits: pandas CSV + labels) is out of the hot path's scope and not rebuilt here."""
from typing import Dict, List

import torch
from torch.nn.functional import pad
from torch.utils.data import default_collate


def collate_fn(xs: List[Dict], packed: bool = False) -> Dict:
    """Zero-pads the ragged per-slide fields to the longest bag and stacks the rest.  Keys: fts [B,N,D], locs [B,N,2],
    ctx_patch [B,N,K,Dp], parent_inds [B,N], num_ims [B] (+ labels, + "slide" objects when present).
    packed=True adds "fts_packed" [sum(num_ims), D]: the same feature rows without the padding, slide after slide -- the
    form utils.inference_end2end ships to the device when it is present (the padded copy then stays on the host)."""
    ragged = {k: [x.pop(k) for x in xs] for k in ("fts", "locs", "ctx_patch", "parent_inds")}
    sizes = [t.shape[0] for t in ragged["locs"]]
    widest = max(sizes)

    def stack_padded(ts, trailing_dims):
        return torch.stack([pad(t, (0, 0) * trailing_dims + (0, widest - t.shape[0])) for t in ts])

    out = {
        "fts": stack_padded(ragged["fts"], 1),
        "locs": stack_padded(ragged["locs"], 1),
        "parent_inds": stack_padded(ragged["parent_inds"], 0),
        "num_ims": torch.tensor(sizes, dtype=torch.long),
    }
    cp0 = ragged["ctx_patch"][0]
    if cp0.shape[1] == 0:        # F.pad rejects zero-sized middle dims (reference dataset.py:221-225)
        out["ctx_patch"] = torch.zeros((len(xs), widest, 0, cp0.shape[2]), dtype=cp0.dtype, device=cp0.device)
    else:
        out["ctx_patch"] = stack_padded(ragged["ctx_patch"], 2)
    if packed:
        out["fts_packed"] = torch.cat(ragged["fts"], 0)
    extra = {"slide": [x.pop("slide") for x in xs]} if "slide" in xs[0] else {}
    return default_collate(xs) | out | extra
