"""Synthetic feature pyramids in the on-disk format of the reference preprocessor.

A slide is a list of L grids, grid l having shape [W0*2^l, H0*2^l, D] float32, where an all-zero
feature vector marks a background cell (reference: preprocess/preprocess.py:89,105 writes zeros for
background; data_utils/slide.py:324 tests `fts[x, y].sum() != 0`).  Definition follows SURVEY.md §8(d):
unit-variance features, every cell independently background with probability p_bg, labels derived
from the slide id.  The generator is seeded per (slide, level) so a subset of a batch can be rebuilt
bit-identically anywhere (CPU generator; tensors are moved to the device afterwards).
"""
from __future__ import annotations

from dataclasses import dataclass
from typing import List, Sequence, Tuple

import torch


@dataclass
class SynthSlide:
    slide_id: int
    grids: List[torch.Tensor]          # L tensors [W_l, H_l, D] float32
    survival_bin: int
    censored: int
    survival: float

    @property
    def num_levels(self) -> int:
        return len(self.grids)


def grid_shape(w0: int, h0: int, level: int) -> Tuple[int, int]:
    return w0 << level, h0 << level


def make_grid(slide_id: int, level: int, w: int, h: int, dim: int, p_bg: float,
              device: torch.device | str = "cpu", seed: int = 1000) -> torch.Tensor:
    """One [w, h, dim] grid.  CPU generator => identical bits wherever it is called."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed + 7919 * slide_id + 104729 * level)
    fts = torch.randn((w, h, dim), generator=g, dtype=torch.float32)
    bg = torch.rand((w, h), generator=g) < p_bg
    fts[bg] = 0.0
    return fts.to(device)


def make_slide(slide_id: int, w0: int, h0: int, num_levels: int, dim: int = 1024, p_bg: float = 0.3,
               device: torch.device | str = "cpu", seed: int = 1000) -> SynthSlide:
    grids = [make_grid(slide_id, l, *grid_shape(w0, h0, l), dim, p_bg, device, seed) for l in range(num_levels)]
    return SynthSlide(slide_id, grids, survival_bin=slide_id % 4, censored=slide_id % 2,
                      survival=10.0 * slide_id + 1.0)


def brca_shape(slide_id: int, seed: int = 1000, wrange=(5, 10), hrange=(4, 8)) -> Tuple[int, int]:
    """Level-0 grid extents of a BRCA-shaped slide: W0~U{5..10}, H0~U{4..8} (SURVEY.md §8(d) cfg 2/3)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed * 31 + slide_id)
    w0 = int(torch.randint(wrange[0], wrange[1] + 1, (1,), generator=g))
    h0 = int(torch.randint(hrange[0], hrange[1] + 1, (1,), generator=g))
    return w0, h0


def make_brca_batch(slide_ids: Sequence[int], num_levels: int = 5, dim: int = 1024, p_bg: float = 0.3,
                    device: torch.device | str = "cpu", seed: int = 1000, wrange=(5, 10), hrange=(4, 8)
                    ) -> List[SynthSlide]:
    return [make_slide(i, *brca_shape(i, seed, wrange, hrange), num_levels, dim, p_bg, device, seed)
            for i in slide_ids]


def make_device_slide_fast(slide_id: int, w0: int, h0: int, num_levels: int, dim: int, p_bg: float,
                           device, seed: int = 1000) -> SynthSlide:
    """Bench-only generator: values drawn with the device RNG (not reproducible on the CPU).  Used for
    the large resident pyramids of bench.py, where only shapes/statistics matter; parity tests use
    make_slide."""
    g = torch.Generator(device=device)
    g.manual_seed(seed + 7919 * slide_id)
    grids = []
    for l in range(num_levels):
        w, h = grid_shape(w0, h0, l)
        fts = torch.randn((w, h, dim), generator=g, dtype=torch.float32, device=device)
        bg = torch.rand((w, h), generator=g, device=device) < p_bg
        fts.mul_((~bg).unsqueeze(-1).to(fts.dtype))
        grids.append(fts)
    return SynthSlide(slide_id, grids, survival_bin=slide_id % 4, censored=slide_id % 2,
                      survival=10.0 * slide_id + 1.0)


def synth_state_dict(shapes: dict, seed: int = 0) -> dict:
    """Deterministic weights for a {key: shape} map (e.g. from model.state_dict()), independent of
    module-construction order: every tensor is drawn from its own generator keyed by its name.
    Matrices ~ U(-1/sqrt(fan_in), 1/sqrt(fan_in)) (nn.Linear's default range), LayerNorm weights
    1 + 0.1 N(0,1), biases/special tokens 0.1 N(0,1)."""
    import zlib
    out = {}
    for key in sorted(shapes):
        shape = tuple(shapes[key])
        g = torch.Generator(device="cpu")
        g.manual_seed((zlib.crc32(key.encode()) + 1000003 * seed) & 0x7FFFFFFF)
        if len(shape) >= 2:
            bound = 1.0 / (shape[-1] ** 0.5)
            t = (torch.rand(shape, generator=g) * 2 - 1) * bound
        elif ".norm" in key and key.endswith("weight"):
            t = 1.0 + 0.1 * torch.randn(shape, generator=g)
        else:
            t = 0.1 * torch.randn(shape, generator=g)
        out[key] = t.float()
    return out
