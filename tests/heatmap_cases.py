"""Synthetic hierarchical patch sets for the heat-map tests, shaped like heatmap_visualise.py builds them: level-0 grid cells,
then the 4 children of the kept patches (some dropped as background), optionally overlapping / off-canvas rectangles."""
import numpy as np


def make_case(seed, L, P, H, W, n0, keep, overlap=False, offgrid=False):
    rng = np.random.default_rng(seed)
    gy, gx = H // P + 1, W // P + 1
    cells = rng.permutation(gy * gx)[:n0]
    locs = [np.stack([(cells // gx) * P, (cells % gx) * P], 1).astype(np.int64)]
    imps = [rng.random(len(cells), dtype=np.float32)]
    for d in range(1, L):
        k = min(keep, len(locs[-1]))
        kept = locs[-1][np.argsort(-imps[-1])[:k]] * 2
        ch = np.concatenate([kept + np.array([dy, dx]) * P for dy in (0, 1) for dx in (0, 1)])
        ch = ch[rng.random(len(ch)) > 0.25]
        if overlap and len(ch) > 2:
            ch = np.concatenate([ch, ch[:2] + P // 2])          # overlapping rectangles: the later patch must win
        locs.append(ch.astype(np.int64)); imps.append(rng.random(len(ch), dtype=np.float32))
    if offgrid:
        locs[0] = np.concatenate([locs[0], np.array([[-P // 2, -P // 3], [H - P // 4, W - P // 5], [H + 5 * P, 0]])])
        imps[0] = np.concatenate([imps[0], rng.random(3, dtype=np.float32)])
    return locs, imps


# the cases whose reference outputs are committed (tests/golden/heatmap_ref.npz, made by tests/golden/make_golden_heatmap.py)
GOLDEN_CASES = [
    dict(seed=10, L=5, P=64, H=450, W=575, n0=30, keep=20),
    dict(seed=11, L=5, P=64, H=277, W=249, n0=12, keep=20, overlap=True),
    dict(seed=12, L=3, P=32, H=250, W=320, n0=40, keep=8, offgrid=True),
    dict(seed=13, L=1, P=32, H=100, W=90, n0=5, keep=1),
    dict(seed=14, L=6, P=128, H=350, W=450, n0=6, keep=32),
]
