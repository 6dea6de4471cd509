"""PreprocessedSlide: the reference's per-slide container (data_utils/slide.py:227-391) re-designed for HBM residency.

The reference keeps every magnification's feature grid in host RAM and, each step and level, gathers the selected
children on the CPU (`iter`, slide.py:277-360).  Here a slide's grids are uploaded ONCE into a PyramidStore (device
memory, reference grid format) together with per-grid occupancy bitmaps; the per-step recursion then runs entirely on
the device (csrc/select.cu) and only names the slide by its store index.  The constructor, `todict()` schema and
`load_patch_preprocessed_slide` signature are the reference's.
"""
import os
from typing import Dict, List, Optional, Sequence, Tuple

import torch

from ..engine import PyramidStore
from ..preprocess import loader

_stores: Dict[tuple, PyramidStore] = {}


def default_store(num_levels: int, dim: int, device=None) -> PyramidStore:
    device = torch.device(device if device is not None else "cuda")
    if device.type == "cuda" and device.index is None:
        device = torch.device("cuda", torch.cuda.current_device())
    key = (num_levels, dim, str(device))
    if key not in _stores:
        _stores[key] = PyramidStore(num_levels, dim, device)
    return _stores[key]


def reset_stores() -> None:
    _stores.clear()


class PreprocessedSlide:
    def __init__(self, path: str, base_power: float, num_levels: int, patch_size: int, ctx_slide: torch.Tensor,
                 ctx_patch_dim: int = None, subtype=None, grids: Optional[Sequence[torch.Tensor]] = None,
                 store: Optional[PyramidStore] = None, device=None):
        self.path = path
        self.patch_size = patch_size
        self.base_power = base_power
        self.ctx_slide = ctx_slide
        self.ctx_patch_dim = ctx_patch_dim
        self.subtype = subtype
        self.slide_id = ".".join(os.path.split(path)[-1].split(".")[:-1])
        if grids is None:                                            # reference slide.py:251-255 (x2 between levels)
            grids = [loader.load(self.slide_id, base_power * 2 ** i) for i in range(num_levels)]
        dim = int(grids[0].shape[-1])
        self.store = store if store is not None else default_store(num_levels, dim, device)
        self.store_index = self.store.add_slide(grids)
        self.grid_shapes = self.store.shapes[self.store_index]
        # level 0 = every grid cell, background included, x-major (reference slide.py:258-271)
        x, y = self.grid_shapes[0]
        ii, jj = torch.meshgrid(torch.arange(x), torch.arange(y), indexing="ij")
        self.locs = torch.stack((ii.reshape(-1), jj.reshape(-1)), dim=1)
        self.ctx_patch = torch.zeros((x * y, 0, ctx_patch_dim))
        self.parent_inds = torch.arange(x * y)
        self.fts: List[torch.Tensor] = [grids[0].reshape(x * y, dim).float().cpu()] + list(self.store.grids[self.store_index][1:])
        self.size_pixels = None

    def load_patches(self, wsi=None):
        assert wsi is None
        return

    def iter(self, *args, **kwargs):
        raise NotImplementedError(
            "paths_b200 performs the level-to-level recursion for the whole batch on the device (paths_engine_select); "
            "use utils.inference_end2end rather than PreprocessedSlide.iter")

    def recurse(self, *args, **kwargs):
        raise NotImplementedError

    def todict(self):
        kwargs = {}
        if self.subtype is not None:
            kwargs["subtype"] = self.subtype
        return {
            "fts": self.fts[0],
            "locs": self.locs * self.patch_size,
            "parent_inds": self.parent_inds,
            "ctx_patch": self.ctx_patch,
            "ctx_slide": self.ctx_slide,
        } | kwargs


def load_patch_preprocessed_slide(path: str, base_power: float, patch_size: int, ctx_dim: Tuple[int, int], num_levels: int,
                                  subtype=None, grids=None, store=None, device=None) -> PreprocessedSlide:
    ctx_slide = torch.zeros((0, ctx_dim[0]))
    return PreprocessedSlide(path, base_power, num_levels, patch_size, ctx_slide, ctx_dim[1], subtype=subtype,
                             grids=grids, store=store, device=device)
