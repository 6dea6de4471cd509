"""CPU checks of the heat-map oracle: bit for bit against outputs of the reference's own compositing code (executed by
tests/golden/make_golden_heatmap.py, committed as tests/golden/heatmap_ref.npz), and against hand-computed answers."""
import os

import numpy as np
import pytest

from tests.heatmap_cases import GOLDEN_CASES, make_case

from oracle import heatmap_oracle as ho


def test_two_level_fold_known_answer():
    # level 0: one 4x4 patch at (0,0) with importance 0.5; level 1: one patch at level-1 pixel (4,4) -> level-0 pixel (2,2), size 2
    out, alpha = ho.composite([np.array([[0, 0]]), np.array([[4, 4]])], [np.array([0.5], np.float32), np.array([0.25], np.float32)], 4, (6, 6))
    a = float(np.float32(0.5) + np.float32(1e-4)); b = float(np.float32(0.25) + np.float32(1e-4))
    want = np.zeros((6, 6)); want[0:4, 0:4] = a; want[2:4, 2:4] = a + 0.5 * b
    assert alpha[0, 0] == 0.5 and alpha[5, 5] == 0 and alpha.sum() == 0.5 * 16
    want[want == 0] = a                                     # zeros take the smallest positive value
    assert np.array_equal(out, want)


def test_later_patch_overwrites_and_clipping():
    out, _ = ho.composite([np.array([[0, 0], [2, 2], [-2, 5]])], [np.array([0.1, 0.9, 0.3], np.float32)], 4, (6, 6))
    hi = float(np.float32(0.9) + np.float32(1e-4)); lo = float(np.float32(0.1) + np.float32(1e-4)); mid = float(np.float32(0.3) + np.float32(1e-4))
    assert out[1, 1] == lo and out[3, 3] == hi and out[5, 5] == hi and out[1, 5] == mid and out[0, 4] == lo
    assert ho.convert_pix(-3, 1, 0) == -2 and ho.convert_pix(256, 3, 0) == 32


@pytest.mark.parametrize("ci", range(len(GOLDEN_CASES)))
def test_oracle_reproduces_the_reference_block(ci):
    ref = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "heatmap_ref.npz"))
    kw = dict(GOLDEN_CASES[ci]); L, P, H, W = kw.pop("L"), kw.pop("P"), kw.pop("H"), kw.pop("W")
    locs, imps = make_case(L=L, P=P, H=H, W=W, **kw)
    out, alpha = ho.composite(locs, imps, P, (H, W))
    assert out.dtype == np.float64 and np.array_equal(out.view(np.int64), ref[f"imp{ci}"].view(np.int64))
    assert np.array_equal(np.asarray(alpha, dtype=np.float64), ref[f"alpha{ci}"])
