"""The reference's per-module methods on the mirror classes -- LSTMCell.forward (model/interface.py:31-58),
TransformerAggregator.forward (model/aggregator.py:58-76), PreprocessedSlide.iter (data_utils/slide.py:277-360) -- each routed to
a single-op entry of the C ABI, against the oracle's restatement of the same method."""
import pytest
import torch

from oracle import paths_oracle as po
from tests.common import (load_fixture, fixture_inputs, edge_grid_edit, rel_err, product_config, product_model, product_batch)

pytestmark = pytest.mark.gpu


@pytest.mark.parametrize("name", ["brca_b4", "small_full"])
def test_lstm_cell_forward(name, monkeypatch):
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    D, C = ocfg.patch_embed_dim, ocfg.hierarchical_ctx_mlp_hidden_dim
    g = torch.Generator().manual_seed(3)
    xs, hs, cs = torch.randn(3, 37, D, generator=g), torch.randn(3, 37, D, generator=g), torch.randn(3, 37, C, generator=g)
    want_h, want_c = po.lstm_cell(sd, xs, hs, cs)
    for prec, tol in (("fp32", 1e-5), ("bf16x3", 1e-4), ("bf16", 2e-2)):
        monkeypatch.setenv("PATHS_B200_PRECISION", prec)
        got_h, got_c = model.lstm(xs.cuda(), hs.cuda(), cs.cuda())
        assert got_h.shape == want_h.shape and got_c.shape == want_c.shape
        assert rel_err(got_h.cpu(), want_h) < tol and rel_err(got_c.cpu(), want_c) < tol, prec
    with pytest.raises(AssertionError):
        model.lstm(xs.cuda()[..., :-1], hs.cuda(), cs.cuda())


@pytest.mark.parametrize("name", ["brca_b4", "small_full"])
def test_transformer_aggregator_forward(name):
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    model.eval()
    model.runtime().set_precision(1)
    d = ocfg.trans_dim
    g = torch.Generator().manual_seed(5)
    num = torch.tensor([5, 1, 17, 9])
    tokens = torch.randn(4, 17, d, generator=g)
    for depth in (0, ocfg.num_levels - 1):
        want = po.aggregate(sd, ocfg, depth, tokens, num)
        got = model.procs[depth].global_agg(torch.zeros(4, 0, d, device="cuda"), tokens.cuda(), None, num.cuda())
        assert got.shape == (4, d) and rel_err(got.cpu(), want) < 1e-4, depth
    with pytest.raises(NotImplementedError):
        model.procs[0].global_agg(torch.zeros(4, 2, d, device="cuda"), tokens.cuda(), num.cuda(), num.cuda())


@pytest.mark.parametrize("name", ["small_edge", "brca_b4"])
def test_preprocessed_slide_iter(name):
    """Level by level, slide by slide, on the oracle's own padded batches: the device select + gather of one slide returns the
    reference's dict (fts, ctx_patch, ctx_slide, locs, parent_inds) -- including the no-child fallback of small_edge."""
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx, edge_grid_edit if name == "small_edge" else None)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    pbatch, store = product_batch(cfg, model, slides)
    batch = po.collate([po.level0_items(s.grids, ocfg) for s in slides])
    saw_fallback = False
    g = torch.Generator().manual_seed(11)
    for depth in range(ocfg.num_levels - 1):
        out = po.level_forward(sd, ocfg, depth, batch)
        # distinct importances: the model's own scores tie exactly on background cells, and torch.topk leaves tie order unspecified
        out["importance"] = torch.rand(out["importance"].shape, generator=g)
        items = []
        for j, s in enumerate(slides):
            n = int(batch["num_ims"][j])
            rec = {}
            want = po.select_expand(s.grids[depth + 1], n, batch["locs"][j], batch["ctx_slide"][j], batch["ctx_patch"][j], out["importance"][j],
                                    out["ctx_slide"][j], out["ctx_patch"][j], ocfg.top_k_patches[depth], ocfg.patch_size, rec)
            saw_fallback |= rec["fallback"]
            got = pbatch["slide"][j].iter(depth, n, batch["locs"][j].clone(), batch["ctx_slide"][j], batch["ctx_patch"][j],
                                          out["importance"][j].detach(), out["ctx_slide"][j].detach(), out["ctx_patch"][j].detach(),
                                          ocfg.top_k_patches[depth])
            assert torch.equal(got["locs"].cpu(), want["locs"]) and torch.equal(got["parent_inds"].cpu(), want["parent_inds"]), (depth, j)
            assert torch.equal(got["fts"].cpu(), want["fts"]), (depth, j)
            assert torch.equal(got["ctx_patch"].cpu(), want["ctx_patch"].detach()) and torch.equal(got["ctx_slide"].cpu(), want["ctx_slide"].detach())
            items.append(want)
        batch = po.collate(items)
    assert saw_fallback == (name == "small_edge")
