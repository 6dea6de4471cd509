"""TEST INFRASTRUCTURE (CPU oracle) -- numpy restatement of the heat-map compositing block of the reference,
heatmap_visualise.py:66-76 (convert_pix / to_pix_space) and 142-176 (rectangle fill, coarse-to-fine fold, zero fill, alpha).

Parity PINNED: the block is inline code of a script function that needs matplotlib / tiatoolbox (absent from this image),
so it cannot be imported -- tests/golden/make_golden_heatmap.py instead reads the reference file, slices the helper
definitions and the block (heatmap_visualise.py:68-76, 142-178) as text and executes them with `Rectangle` / `ax` stubbed;
its outputs are committed (tests/golden/heatmap_ref.npz) and tests/test_heatmap_oracle.py holds this restatement to them
bit for bit.  The GPU kernel is held to both (tests/test_gpu_heatmap.py).  Only tests/ may import this module.

Note on dtype: `imp[i] + 1e-4` adds a python float to a numpy float32 scalar; under this image's numpy 2 (NEP 50) that is a
float32 add (under the reference's pinned numpy 1.x it would have been float64).  The value is then stored in the float64
canvas.  The kernel follows numpy 2, the behaviour a user of this image gets.
"""
import numpy as np


def convert_pix(pix, depth, to_depth):                       # heatmap_visualise.py:67-72
    e = to_depth - depth
    if e <= 0:
        return pix // 2 ** (-e)
    return pix * 2 ** e


def composite(locs_per_depth, imps, P, shape):
    L = len(imps)
    s1, s2 = shape
    overall_imp = np.zeros((L,) + tuple(shape))              # :142
    for depth in range(L):                                   # :147-165
        locs = np.asarray(locs_per_depth[depth])
        size = convert_pix(P, depth, 0)
        imp = np.asarray(imps[depth], dtype=np.float32)
        for i in range(locs.shape[0]):
            y, x = (int(v) for v in locs[i].tolist())
            y, x = convert_pix(y, depth, 0), convert_pix(x, depth, 0)
            overlap_y = max(y, 0) <= min(y + size, s1)
            overlap_x = max(x, 0) <= min(x + size, s2)
            if overlap_y and overlap_x:
                y1, y2 = max(y, 0), min(y + size, s1)
                x1, x2 = max(x, 0), min(x + size, s2)
                overall_imp[depth, y1:y2, x1:x2] = imp[i] + 1e-4
    for depth in range(L - 2, -1, -1):                       # :168-170
        relevant_mask = overall_imp[depth + 1] != 0
        overall_imp[depth][relevant_mask] = overall_imp[depth][relevant_mask] + overall_imp[depth + 1][relevant_mask] * 0.5
    overall_imp = overall_imp[0]                             # :172
    alpha = np.where(overall_imp > 0, 0.5, 0)                # :174
    if (overall_imp > 0).any():
        overall_imp[overall_imp == 0] = np.min(overall_imp[overall_imp > 0])   # :175
    return overall_imp, alpha
