"""CPU checks of the heat-map oracle against hand-computed answers (the reference block cannot be imported: matplotlib)."""
import numpy as np

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
