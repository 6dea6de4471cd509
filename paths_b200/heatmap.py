"""Importance heat-map compositing on the GPU: the numpy block of heatmap_visualise.py:142-178 as one call.

    overall_imp, alpha = composite_heatmap(locs_per_depth, imps, patch_size, (H, W))

`locs_per_depth[d]` is the `[n_d, 2]` (y, x) integer tensor `slide_depths[d].locs` of the reference loop
(heatmap_visualise.py:146), `imps[d]` the `[n_d]` float32 importances it collected (heatmap_visualise.py:100).  Returns the
float64 `[H, W]` map handed to `ax.imshow` and its alpha mask; plotting itself (matplotlib) stays with the caller.
There is no CPU fallback: the kernels live in libpaths_b200.so.
"""
from __future__ import annotations

import torch

from . import _lib


def composite_heatmap(locs_per_depth, imps, patch_size: int, shape, device=None):
    assert len(locs_per_depth) == len(imps) and len(imps) >= 1, "one locs tensor and one importance vector per level"
    H, W = int(shape[0]), int(shape[1])
    L = len(imps)
    device = torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())
    assert device.type == "cuda", "composite_heatmap runs on a CUDA device (no CPU fallback)"
    locs, vals, off = [], [], [0]
    for lc, im in zip(locs_per_depth, imps):
        lc = torch.as_tensor(lc).reshape(-1, 2)
        im = torch.as_tensor(im).reshape(-1)
        assert lc.shape[0] == im.shape[0], f"locs {tuple(lc.shape)} and importance {tuple(im.shape)} disagree"
        locs.append(lc.to(torch.int32)); vals.append(im.to(torch.float32)); off.append(off[-1] + lc.shape[0])
    locs_d = torch.cat(locs).contiguous().to(device)
    vals_d = torch.cat(vals).contiguous().to(device)
    off_d = torch.tensor(off, dtype=torch.int32, device=device)
    lib = _lib.lib()
    wsb = lib.paths_heatmap_workspace_bytes(L, H, W)
    ws = torch.empty(wsb, dtype=torch.uint8, device=device)
    out = torch.empty((H, W), dtype=torch.float64, device=device)
    alpha = torch.empty((H, W), dtype=torch.float32, device=device)
    st = torch.cuda.current_stream(device).cuda_stream
    _lib.check(lib.paths_heatmap_composite(locs_d.data_ptr(), vals_d.data_ptr(), off_d.data_ptr(), off[-1], L, int(patch_size), H, W,
                                           ws.data_ptr(), wsb, out.data_ptr(), alpha.data_ptr(), st), "heatmap_composite")
    return out, alpha
