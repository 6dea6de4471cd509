"""Heat-map compositing: CUDA kernels vs the numpy restatement of heatmap_visualise.py:142-176, bit-exact."""
import numpy as np
import pytest
import torch

from oracle import heatmap_oracle as ho
from tests.heatmap_cases import GOLDEN_CASES, make_case

pytestmark = pytest.mark.gpu


_case = make_case


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


@pytest.mark.parametrize("ci", range(len(GOLDEN_CASES)))
def test_heatmap_matches_the_reference_outputs(ci):
    """The CUDA compositing against outputs of the reference's own code (tests/golden/heatmap_ref.npz), bit for bit."""
    import os
    from paths_b200.heatmap import composite_heatmap
    ref = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "heatmap_ref.npz"))
    kw = dict(GOLDEN_CASES[ci]); L, P, H, W = kw.pop("L"), kw.pop("P"), kw.pop("H"), kw.pop("W")
    locs, imps = make_case(L=L, P=P, H=H, W=W, **kw)
    got, alpha = composite_heatmap([torch.from_numpy(l) for l in locs], [torch.from_numpy(i) for i in imps], P, (H, W), device="cuda:0")
    assert np.array_equal(got.cpu().numpy().view(np.int64), ref[f"imp{ci}"].view(np.int64))
    assert np.array_equal(alpha.cpu().numpy().astype(np.float64), ref[f"alpha{ci}"])
