"""PatchBatch: the padded single-magnification batch object of the reference API (data_utils/patch_batch.py:13-75)."""
from typing import Dict

import torch


def _todevice(x, device):
    if hasattr(x, "to"):
        return x.to(device)
    if isinstance(x, (list, tuple)):
        return [_todevice(i, device) for i in x]
    return x


class PatchBatch:
    """locs [B,N,2] (pixel coordinates at this magnification), num_ims [B], parent_inds [B,N], ctx_slide [B,depth,Ds],
    ctx_patch [B,N,depth,Dp], fts [B,N,D].  Same shape/device checks and `valid_inds` mask as the reference."""

    def __init__(self, locs, num_ims, parent_inds, ctx_slide, ctx_patch, fts, **unused_kwargs):
        batch_size, max_patches, _ = fts.shape
        _, self.ctx_depth, self.ctx_dim1 = ctx_slide.shape
        self.ctx_dim2 = ctx_patch.shape[-1]
        assert locs.shape == (batch_size, max_patches, 2)
        assert num_ims.shape == (batch_size,)
        assert parent_inds.shape == (batch_size, max_patches)
        assert ctx_slide.shape == (batch_size, self.ctx_depth, self.ctx_dim1)
        assert ctx_patch.shape == (batch_size, max_patches, self.ctx_depth, self.ctx_dim2)
        assert num_ims.max().item() == max_patches
        self.device = fts.device
        assert self.device == locs.device == num_ims.device == parent_inds.device == ctx_slide.device == ctx_patch.device
        self.batch_size, self.max_patches = batch_size, max_patches
        self.fts, self.locs, self.num_ims = fts, locs, num_ims
        self.parent_inds, self.ctx_slide, self.ctx_patch = parent_inds, ctx_slide, ctx_patch
        self.valid_inds = torch.arange(max_patches, device=num_ims.device).expand(batch_size, -1) < num_ims[:, None]


def from_batch(batch: Dict, device) -> PatchBatch:
    return PatchBatch(**{k: _todevice(v, device) for k, v in batch.items()})
