"""Heat-map compositing: CUDA kernels vs the numpy restatement of heatmap_visualise.py:142-176, bit-exact."""
import numpy as np
import pytest
import torch

from oracle import heatmap_oracle as ho

pytestmark = pytest.mark.gpu


def _case(seed, L, P, H, W, n0, keep, overlap=False, offgrid=False):
    """Hierarchical patch sets like heatmap_visualise.py builds them: level-0 grid cells, then 4 children of the kept patches."""
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


@pytest.mark.parametrize("cfg", [
    dict(seed=0, L=5, P=256, H=1800, W=2300, n0=30, keep=20),
    dict(seed=1, L=5, P=256, H=1111, W=997, n0=12, keep=20, overlap=True),
    dict(seed=2, L=3, P=64, H=500, W=640, n0=40, keep=8, offgrid=True),
    dict(seed=3, L=1, P=32, H=100, W=90, n0=5, keep=1),
    dict(seed=4, L=6, P=256, H=700, W=900, n0=6, keep=32),
])
def test_heatmap_matches_oracle_bit_exact(cfg):
    from paths_b200.heatmap import composite_heatmap
    kw = dict(cfg); L, P, H, W = kw.pop("L"), kw.pop("P"), kw.pop("H"), kw.pop("W")
    locs, imps = _case(L=L, P=P, H=H, W=W, **kw)
    want, want_alpha = ho.composite(locs, imps, P, (H, W))
    got, alpha = composite_heatmap([torch.from_numpy(l) for l in locs], [torch.from_numpy(i) for i in imps], P, (H, W), device="cuda:0")
    got, alpha = got.cpu().numpy(), alpha.cpu().numpy()
    assert got.dtype == np.float64 and got.shape == (H, W)
    assert np.array_equal(got.view(np.int64), want.view(np.int64)), f"max |diff| {np.abs(got - want).max()}"
    assert np.array_equal(alpha.astype(np.float64), want_alpha.astype(np.float64))


def test_heatmap_empty_levels():
    from paths_b200.heatmap import composite_heatmap
    locs = [np.zeros((0, 2), np.int64), np.zeros((0, 2), np.int64)]
    imps = [np.zeros(0, np.float32), np.zeros(0, np.float32)]
    want, want_alpha = ho.composite(locs, imps, 256, (64, 48))
    got, alpha = composite_heatmap([torch.from_numpy(l) for l in locs], [torch.from_numpy(i) for i in imps], 256, (64, 48), device="cuda:0")
    assert np.array_equal(got.cpu().numpy(), want) and np.array_equal(alpha.cpu().numpy().astype(np.float64), want_alpha.astype(np.float64))
