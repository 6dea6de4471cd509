"""Pins oracle/heatmap_oracle.py to the reference by EXECUTING the reference's own compositing code.

    python tests/golden/make_golden_heatmap.py        (build container only: reads /root/reference)

heatmap_visualise.py cannot be imported here (matplotlib, tiatoolbox and sksurv are absent and the block is inline code of a
script function), so this script reads the file, takes the two helper definitions (`convert_pix`, `to_pix_space`) and the
compositing block -- from `shape = (H, W)` to the zero-fill line, reference lines 142-175 -- as TEXT at run time, and
`exec`s them with the two matplotlib objects the block touches (`Rectangle`, `ax.add_patch`) stubbed.  Nothing of the
reference is written into the repository; the outputs for the cases of tests/heatmap_cases.py go to
tests/golden/heatmap_ref.npz and tests/test_heatmap_oracle.py holds the oracle to them bit for bit.
"""
import os
import sys
import textwrap

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PATHS_REFERENCE", "/root/reference")
sys.path.insert(0, ROOT)
from tests.heatmap_cases import GOLDEN_CASES, make_case  # noqa: E402


def slice_between(lines, first_marker, last_marker):
    a = next(i for i, l in enumerate(lines) if first_marker in l)
    b = next(i for i, l in enumerate(lines) if last_marker in l and i >= a)
    return a, b, textwrap.dedent("".join(lines[a:b + 1]))


def main():
    lines = open(os.path.join(REF, "heatmap_visualise.py")).readlines()
    h0, h1, helpers = slice_between(lines, "def convert_pix(pix, depth, to_depth):", "return convert_pix(y, depth, 0), convert_pix(x, depth, 0)")
    b0, b1, block = slice_between(lines, "shape = (H, W)", "overall_imp[overall_imp == 0] = np.min(overall_imp[overall_imp > 0])")
    print(f"helpers = heatmap_visualise.py:{h0 + 1}-{h1 + 1}, block = heatmap_visualise.py:{b0 + 1}-{b1 + 1}")

    class _Ax:
        def add_patch(self, r):
            pass

    class _Slide:
        def __init__(self, locs):
            self.locs = torch.from_numpy(locs)              # the reference holds patch origins as a LongTensor (slide.py)

    out = {}
    for ci, cfg in enumerate(GOLDEN_CASES):
        kw = dict(cfg); L, P, H, W = kw.pop("L"), kw.pop("P"), kw.pop("H"), kw.pop("W")
        locs, imps = make_case(L=L, P=P, H=H, W=W, **kw)
        ns = {"np": np, "Rectangle": lambda *a, **k: None, "ax": _Ax(), "L": L, "P": P, "H": H, "W": W,
              "slide_depths": [_Slide(l) for l in locs], "imps": [i.copy() for i in imps]}
        exec(helpers, ns)
        exec(block, ns)
        out[f"imp{ci}"] = ns["overall_imp"]
        out[f"alpha{ci}"] = np.asarray(ns["alpha"], dtype=np.float64)
        print(ci, cfg, "covered", int((ns["alpha"] > 0).sum()), "min", float(ns["overall_imp"].min()), "max", float(ns["overall_imp"].max()))
    out["numpy"] = np.array(np.__version__)
    np.savez_compressed(os.path.join(HERE, "heatmap_ref.npz"), **out)


if __name__ == "__main__":
    main()
