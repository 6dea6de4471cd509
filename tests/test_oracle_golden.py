"""Pins oracle/paths_oracle.py to the reference: every committed fixture (made by running the unmodified
reference, tests/golden/make_golden.py) must be reproduced by the oracle restatement."""
import pytest
import torch

from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, edge_grid_edit, rel_err

CASES = ["cfg1_sample", "brca_b4", "small_full", "small_edge", "small_rnn", "small_concat1d"]
TOL = 2e-5   # fp32 vs fp32, different op order only


@pytest.mark.parametrize("name", CASES)
def test_oracle_matches_reference(name):
    fx = load_fixture(name)
    cfg, sd, slides, ybin, cens = fixture_inputs(fx, edge_grid_edit if name == "small_edge" else None)
    hz, loss, grads, recs = po.grads(sd, cfg, [s.grids for s in slides], ybin, cens)
    assert len(recs) == len(fx["levels"])
    for l, (rec, ref) in enumerate(zip(recs, fx["levels"])):
        assert torch.equal(rec.num_ims, ref["num_ims"]), f"level {l} bag sizes"
        assert torch.equal(rec.locs, ref["locs"]), f"level {l} patch locations (selection/expansion)"
        assert rel_err(rec.importance, ref["importance"]) < TOL
        assert rel_err(rec.logits, ref["logits"]) < TOL
        assert rel_err(rec.ctx_slide, ref["ctx_slide"]) < TOL
        assert rel_err(rec.ctx_patch[:, :4], ref["ctx_patch_head"]) < TOL
    assert rel_err(hz, fx["hazards"]) < TOL
    assert abs(float(loss) - float(fx["loss"])) < TOL * max(1.0, abs(float(fx["loss"])))
    gmax = max(g["vals"].abs().max().item() for g in fx["grads"].values() if g is not None)
    for k, g in fx["grads"].items():
        if g is None:
            assert grads[k] is None or float(grads[k].abs().max()) == 0.0, k
            continue
        mine = grads[k]
        if mine is None:   # reference gives exact zeros (dead encoder / cross-attn weights)
            assert g["norm"] == 0.0, k
            continue
        if g["norm"] == 0.0:
            assert float(mine.abs().max()) == 0.0, k
            continue
        # error relative to the tensor's own scale, floored at 1e-3 of the largest gradient entry seen
        # (one-element gradients such as importance_mlp.2.bias are cancelling sums of many terms)
        scale = max(float(mine.abs().max()), 1e-3 * gmax)
        assert float((mine.reshape(-1)[g["idx"]] - g["vals"]).abs().max()) / scale < 5 * TOL, k
        assert abs(float(mine.double().norm()) - g["norm"]) / max(g["norm"], 1e-3 * gmax) < 5 * TOL, k
        if "full" in g:
            assert float((mine - g["full"]).abs().max()) / scale < 5 * TOL, k


def test_oracle_fp64_close_to_fp32():
    fx = load_fixture("small_full")
    cfg, sd, slides, ybin, cens = fixture_inputs(fx)
    hz64, loss64, _ = po.end2end(sd, cfg, [[g.double() for g in s.grids] for s in slides], ybin, cens,
                                 dtype=torch.float64)
    assert rel_err(hz64.float(), fx["hazards"]) < 1e-4
