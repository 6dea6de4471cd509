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
        g0 = grids[0].reshape(x * y, dim)
        if self.store.dtype == torch.bfloat16:                       # the batch's level-0 rows are the store's (rounded) rows
            g0 = g0.to(torch.bfloat16)
        self.fts: List[torch.Tensor] = [g0.float().cpu()] + list(self.store.grids[self.store_index][1:])
        self.size_pixels = None

    def load_patches(self, wsi=None):
        assert wsi is None
        return

    def iter(self, magnification_index: int, npatches: int, locs, ctx_slide, ctx_patch, importance, new_ctx_slide,
             new_ctx_patch, keep_patches: int = -1, imp_cpu=None):
        """Data of magnification level `index` -> `index + 1` for THIS slide (reference slide.py:277-360): top-K by importance,
        4-child expansion, bounds + background filter, feature gather, no-child fallback -- one call of the stand-alone device
        entries `paths_select_expand` + `paths_gather_rows` with B = 1 (utils.inference_end2end does the same for the whole batch
        without leaving the device).  Returns the reference's dict; like the reference it floor-divides `locs` in place."""
        import ctypes as C
        from .. import _lib as L
        lib = L.lib()
        dev = self.store.device
        i = magnification_index
        n = int(npatches)
        locs //= self.patch_size                                            # (in place on the caller's tensor, slide.py:280)
        imp = (imp_cpu if imp_cpu is not None else importance)[:n].detach().to(dev, torch.float32).contiguous()
        ctx_patch = ctx_patch[:n]; new_ctx_patch = new_ctx_patch[:n]
        ctx_slide = torch.cat((ctx_slide, new_ctx_slide[None]), dim=0)
        hist = torch.cat((ctx_patch, new_ctx_patch[:, None]), dim=1)        # N x (K+1) x D
        gw, gh = self.grid_shapes[i + 1]
        cells = gw * gh
        k = n if keep_patches == -1 else min(n, int(keep_patches))
        kcap = max(k, 1)
        i32 = dict(dtype=torch.int32, device=dev)
        seg = torch.tensor([0, n], **i32); count = torch.tensor([n], **i32)
        row_xy = locs[:n].to(dev, torch.int32).contiguous()
        sel = torch.empty(kcap, **i32); sel_count = torch.empty(1, **i32); child_of = torch.empty(max(n, 1) * 4, **i32)
        n_count = torch.zeros(1, **i32); n_ntok = torch.zeros(1, **i32); n_seg = torch.zeros(2, **i32)
        n_row_slide = torch.empty(cells, **i32); n_row_xy = torch.empty(cells, 2, **i32)
        n_row_parent = torch.empty(cells, **i32); n_parent_local = torch.empty(cells, **i32)
        overflow = torch.zeros(2, **i32); scratch = torch.zeros(2, **i32)
        sid = torch.tensor([self.store_index], **i32)
        _, st = self.store.tables()
        stream = torch.cuda.current_stream(dev).cuda_stream
        L.check(lib.paths_select_expand(imp.data_ptr(), seg.data_ptr(), row_xy.data_ptr(), count.data_ptr(), 1, max(n, 1), int(keep_patches), kcap,
                                        max(n, 1), C.byref(st), i, self.store.L, sid.data_ptr(), sel.data_ptr(), sel_count.data_ptr(),
                                        child_of.data_ptr(), n_count.data_ptr(), n_ntok.data_ptr(), n_seg.data_ptr(), n_row_slide.data_ptr(),
                                        n_row_xy.data_ptr(), n_row_parent.data_ptr(), n_parent_local.data_ptr(), cells, cells,
                                        overflow.data_ptr(), scratch.data_ptr(), stream), "select_expand")
        m = int(n_count.item())
        assert int(overflow[0].item()) == 0
        new_fts = torch.empty(m, self.store.D, device=dev)
        L.check(lib.paths_gather_rows(C.byref(st), i + 1, self.store.L, sid.data_ptr(), n_row_slide.data_ptr(), n_row_xy.data_ptr(),
                                      new_fts.data_ptr(), self.store.D, m, None, stream), "gather_rows")
        parent_row = n_row_parent[:m].long()
        out_dev = hist.device
        if m and int(parent_row[0].item()) < 0:                             # fallback (slide.py:336-352): zero state, cell indices
            new_ctx = torch.zeros((m, hist.size(1), hist.size(2)), device=out_dev)
        else:
            new_ctx = hist[parent_row.to(out_dev)]                          # a child inherits its parent's whole history (slide.py:318)
        return {
            "fts": new_fts,
            "ctx_patch": new_ctx,
            "ctx_slide": ctx_slide,
            "locs": n_row_xy[:m].to(locs.device, locs.dtype) * self.patch_size,
            "parent_inds": n_parent_local[:m].to(locs.device, locs.dtype),
        }

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
