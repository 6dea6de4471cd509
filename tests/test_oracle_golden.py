"""Pins oracle/paths_oracle.py to the reference: every committed fixture (made by running the unmodified
reference, tests/golden/make_golden.py) must be reproduced by the oracle restatement."""
import pytest
import torch

from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, edge_grid_edit, rel_err

CASES = ["cfg1_sample", "brca_b4", "small_full", "small_edge", "small_rnn", "small_concat1d", "small_subtype"]
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


def test_oracle_dropout_sites_match_torch():
    """Training-mode dropout: the masks torch drew inside the reference aggregator (tests/golden/make_golden_dropout.py)
    replayed through the oracle's hand-written decoder give the module's output and every gradient -- pins where the five
    dropout sites act, that each block is dropped including its bias, and the 1/(1-p) rescaling."""
    import os
    fx = torch.load(os.path.join(os.path.dirname(__file__), "golden", "agg_dropout.pt"), weights_only=False)
    dm = fx["dims"]
    pre = "procs.0.global_agg."
    leaf = {pre + k: v.clone().requires_grad_(True) for k, v in fx["state_dict"].items()}
    tokens = fx["tokens"].clone().requires_grad_(True)
    masks = {(0, j, site): m for (j, site), m in fx["masks"].items()}
    cfg = po.OracleConfig(trans_dim=dm["d"], trans_heads=dm["heads"], trans_layers=dm["layers"])
    out = po.aggregate(leaf, cfg, 0, tokens, fx["num_ims"], drop=po.RecordedDropout(masks))
    assert rel_err(out.detach(), fx["out"]) < TOL
    (out * fx["w"]).sum().backward()
    assert rel_err(tokens.grad, fx["dtokens"]) < 5 * TOL
    live = 0
    for k, g in fx["grads"].items():
        mine = leaf[pre + k].grad
        if g is None or float(g.abs().max()) == 0.0:              # encoder / cross-attention weights: dead
            assert mine is None or float(mine.abs().max()) == 0.0, k
            continue
        assert mine is not None and rel_err(mine, g) < 5 * TOL, k
        live += 1
    assert live >= 30
    # without the masks the same weights give a different answer (the fixture really exercised dropout)
    plain = po.aggregate({k: v.detach() for k, v in leaf.items()}, cfg, 0, fx["tokens"], fx["num_ims"])
    assert rel_err(plain, fx["out"]) > 1e-2


def test_hash_dropout_statistics_and_determinism():
    """The CUDA path's keep/drop convention (paths_b200/csrc/common.cuh: drop_keep) as restated by HashDropout: keep rate
    1-p, survivors scaled by 1/(1-p), decisions a pure function of (seed, level, layer, site, row, column)."""
    rows = torch.arange(4096).view(64, 64)
    for p in (0.05, 0.3):
        d = po.HashDropout(p, seed=99)
        m = d.elem(2, 1, po.DROP_FFN_INNER, rows, 128)
        assert abs(float((m == 0).float().mean()) - p) < 4e-3
        assert torch.all((m == 0) | (m == float(d.inv_keep)))
        assert torch.equal(m, po.HashDropout(p, seed=99).elem(2, 1, po.DROP_FFN_INNER, rows, 128))
        for other in (d.elem(2, 0, po.DROP_FFN_INNER, rows, 128), d.elem(1, 1, po.DROP_FFN_INNER, rows, 128),
                      d.elem(2, 1, po.DROP_FFN_OUT, rows, 128), po.HashDropout(p, seed=100).elem(2, 1, po.DROP_FFN_INNER, rows, 128)):
            agree = float(((m == 0) == (other == 0)).float().mean())
            assert abs(agree - (p * p + (1 - p) ** 2)) < 6e-3          # independent masks
        a = d.attn(0, 0, rows[:4], 4)
        assert a.shape == (4, 4, 64, 64) and abs(float((a == 0).float().mean()) - p) < 8e-3
    assert float((po.HashDropout(0.0, 1).elem(0, 0, 1, rows, 8) == 1).float().mean()) == 1.0
