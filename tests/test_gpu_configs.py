"""Parity on the other BASELINE.json configurations: cfg 2 (brca_paths_0 inference, 64 slides, per-level
importance), cfg 4 (one 20 k-patch bag, K=64, mixed into a varlen batch) and cfg 5 (6 magnifications, K=32).  Sizes
the oracle finishes in seconds are compared value-by-value; the full-size 20 k bag is checked through size-independent
properties (batch invariance, top-K validity and ordering, child geometry, occupancy)."""
import copy

import pytest
import torch

from oracle import paths_oracle as po
from paths_b200 import synth
from tests.common import BASE_CONFIG, product_model, product_batch, rel_err, load_fixture, fixture_inputs, product_config
from tests.parity_check import check_against_oracle

pytestmark = pytest.mark.gpu


def _config(**over):
    from paths_b200.config import Config
    data = copy.deepcopy(BASE_CONFIG)
    for k, v in over.items():
        if k in data["model_config"]:
            data["model_config"][k] = v
        else:
            data[k] = v
    data["model_config"]["dropout"] = 0.0
    return Config.from_dict(data, test_mode=True)


def _weights(cfg):
    shapes = {k: tuple(v.shape) for k, v in cfg.get_model().state_dict().items()}
    return synth.synth_state_dict(shapes, seed=3)


def _labels(slides):
    return (torch.tensor([s.survival_bin for s in slides], dtype=torch.long), torch.tensor([s.censored for s in slides], dtype=torch.long))


def _run(cfg, sd, slides, prec, grad=True):
    from paths_b200.utils import inference_end2end
    model = product_model(cfg, sd)
    rt = model.runtime()
    rt.set_precision(prec)
    batch, store = product_batch(cfg, model, slides)
    if grad:
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
    else:
        model.eval()
        with torch.no_grad():
            hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    torch.cuda.synchronize()
    return model, rt, hz, loss


@pytest.mark.parametrize("prec", [1, 2])
def test_cfg2_inference_importance_64_slides(prec):
    """models/brca_paths_0, batch of 64 synthetic slides, no_grad: per-level importance scores (heatmap_visualise.py:99-100)."""
    cfg = _config()
    sd = _weights(cfg)
    slides = synth.make_brca_batch(range(64), cfg.num_levels, cfg.model_config.patch_embed_dim)
    ybin, cens = _labels(slides)
    model, rt, hz, loss = _run(cfg, sd, slides, prec, grad=False)
    assert all(p.grad is None for p in model.parameters())
    ocfg = po.OracleConfig.from_config(cfg)
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec, check_grads=False)


@pytest.mark.parametrize("prec", [0, 1])
def test_cfg5_six_magnifications_k32(prec):
    """Deep hierarchy: 6 levels (0.3125x-10x), K=32, fwd+bwd."""
    cfg = _config(num_levels=6, base_power=0.3125, top_k_patches=[32] * 5)
    sd = _weights(cfg)
    slides = synth.make_brca_batch(range(4), 6, cfg.model_config.patch_embed_dim, wrange=(3, 5), hrange=(2, 4))
    ybin, cens = _labels(slides)
    model, rt, hz, loss = _run(cfg, sd, slides, prec)
    check_against_oracle(po.OracleConfig.from_config(cfg), sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec)


@pytest.mark.parametrize("prec", [0, 1])
def test_cfg4_large_bag_midsize_against_oracle(prec):
    """One 36x36 = 1296-patch bag (K=64 -> up to 256 children) mixed with BRCA-shaped bags, 3 levels, fwd+bwd."""
    cfg = _config(num_levels=3, top_k_patches=[64, 64])
    sd = _weights(cfg)
    D = cfg.model_config.patch_embed_dim
    slides = [synth.make_slide(100, 36, 36, 3, D)] + synth.make_brca_batch(range(3), 3, D)
    ybin, cens = _labels(slides)
    model, rt, hz, loss = _run(cfg, sd, slides, prec)
    check_against_oracle(po.OracleConfig.from_config(cfg), sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec)


def test_cfg4_20k_patch_bag_properties():
    """Full size: a 142x141 = 20 022-patch level-0 bag, K=64, mixed with 63 BRCA-shaped slides in one varlen batch."""
    from paths_b200.utils import inference_end2end
    cfg = _config(num_levels=3, top_k_patches=[64, 64])
    sd = _weights(cfg)
    D = cfg.model_config.patch_embed_dim
    dev = "cuda"
    big = synth.make_device_slide_fast(500, 142, 141, 3, D, 0.3, dev)
    small = [synth.make_device_slide_fast(i, *synth.brca_shape(i), 3, D, 0.3, dev) for i in range(63)]
    model = product_model(cfg, sd)
    rt = model.runtime()
    rt.set_precision(1)

    def run(slides):
        batch, store = product_batch(cfg, model, slides)
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
        torch.cuda.synchronize()
        eng = rt.last_engine
        out = {"hz": hz.detach().clone(), "loss": float(loss)}
        B = len(slides)
        for l in range(3):
            M = int(eng.buffer(l, "COUNT").item())
            seg = eng.buffer(l, "SEG")[:B + 1].tolist()
            out[l] = {"seg": seg, "alpha": eng.buffer(l, "ALPHA")[:M].clone(), "xy": eng.buffer(l, "ROW_XY").view(-1, 2)[:M].clone(),
                      "parent": eng.buffer(l, "ROW_PARENT")[:M].clone(), "X": eng.buffer(l, "X").view(-1, D)[:M].clone()}
            if l < 2:
                kcap = eng.buffer(l, "SEL").numel() // B
                out[l]["sel"] = eng.buffer(l, "SEL").view(B, kcap).clone()
                out[l]["sel_count"] = eng.buffer(l, "SEL_COUNT")[:B].clone()
        assert not eng.overflowed()[0]
        return out

    mixed = run([big] + small)
    g_mixed = {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}
    assert all(torch.isfinite(g).all() for g in g_mixed.values()) and mixed["loss"] == mixed["loss"]
    seg0 = mixed[0]["seg"]
    assert seg0[1] - seg0[0] == 142 * 141
    # ---- selection: exactly the 64 most important patches, importance-descending (slide.py:294-301)
    a = mixed[0]["alpha"][seg0[0]:seg0[1]]
    sel = mixed[0]["sel"][0, :int(mixed[0]["sel_count"][0])].long() - seg0[0]
    assert len(sel) == 64 and len(set(sel.tolist())) == 64
    s_a = a[sel]
    assert bool((s_a[:-1] >= s_a[1:]).all())
    rest = torch.ones_like(a, dtype=torch.bool); rest[sel] = False
    assert float(s_a.min()) >= float(a[rest].max())
    # ---- children: 4 blocks (2x,2y),(2x,2y+1),(2x+1,2y),(2x+1,2y+1) of the kept parents, in bounds, non-background
    seg1 = mixed[1]["seg"]
    xy1 = mixed[1]["xy"][seg1[0]:seg1[1]]
    par = mixed[1]["parent"][seg1[0]:seg1[1]].long()
    pxy = mixed[0]["xy"][par]
    off = xy1 - 2 * pxy
    assert bool(((off >= 0) & (off <= 1)).all()) and 0 < len(xy1) <= 256
    assert set(par.tolist()) <= set((sel + seg0[0]).tolist())
    blk = (off[:, 0] * 2 + off[:, 1])
    assert bool((blk[:-1] <= blk[1:]).all()), "children must come in the reference's four-block order"
    W1, H1 = big.grids[1].shape[:2]
    assert bool((xy1[:, 0] < W1).all()) and bool((xy1[:, 1] < H1).all())
    feats = big.grids[1][xy1[:, 0].long(), xy1[:, 1].long()]
    assert torch.equal(feats, mixed[1]["X"][seg1[0]:seg1[1]]), "gathered child features must be the grid's rows, bit-exact"
    assert bool((feats.abs().sum(1) != 0).all()), "background children must have been filtered out"
    # ---- batch invariance: the big slide alone gives the same scores and hazards as inside the mixed varlen batch
    alone = run([big])
    n0 = 142 * 141
    assert rel_err(alone[0]["alpha"][:n0], a) < 1e-6
    assert torch.equal(alone[1]["xy"], xy1)
    assert rel_err(alone["hz"][0], mixed["hz"][0]) < 1e-5


@pytest.mark.parametrize("prec", [1, 2])
def test_level_sets_and_operand_paths_agree(prec, monkeypatch):
    """The aggregators of any grouping of the levels (one grouped pass, the default split, one pass per level) and both operand
    paths of their contractions (pre-split planes over TMA / in-kernel conversion) compute the same step: hazards bit-identical
    or within float re-association, gradients within 1e-5 of each other (they differ only in summation order)."""
    from paths_b200.utils import inference_end2end
    fx = load_fixture("brca_b4")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    results = []
    for env in ({}, {"PATHS_B200_DEC_SETS_FWD": "01234", "PATHS_B200_DEC_SETS_BWD": "01234"},
                {"PATHS_B200_DEC_SETS_FWD": "0,1,2,3,4", "PATHS_B200_DEC_SETS_BWD": "0,1,2,3,4"},
                {"PATHS_B200_DEC_SETS_FWD": "01,234", "PATHS_B200_DEC_SETS_BWD": "0,1234", "PATHS_B200_DEC_NO_PLANES": "1"},
                # backward sets inside the forward sets reuse the forward pass's operand planes; sets that straddle them re-split
                {"PATHS_B200_DEC_SETS_FWD": "01234", "PATHS_B200_DEC_SETS_BWD": "0,12,34"},
                {"PATHS_B200_DEC_SETS_FWD": "01,234", "PATHS_B200_DEC_SETS_BWD": "0,1234"},
                {"PATHS_B200_DEC_NO_PLANES": "1"}):
        for k in ("PATHS_B200_DEC_SETS_FWD", "PATHS_B200_DEC_SETS_BWD", "PATHS_B200_DEC_NO_PLANES"):
            monkeypatch.delenv(k, raising=False)
        for k, v in env.items():
            monkeypatch.setenv(k, v)                        # read when the engine is created
        cfg = product_config(fx)
        model = product_model(cfg, sd)
        model.runtime().set_precision(prec)
        batch, _ = product_batch(cfg, model, slides)
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
        torch.cuda.synchronize()
        results.append((hz.detach().clone(), {n: p.grad.detach().clone() for n, p in model.named_parameters() if p.grad is not None}))
    hz0, g0 = results[0]
    tol = 1e-5 if prec == 1 else 2e-3                       # bf16 mode: the two operand paths round at different points
    for hz, g in results[1:]:
        assert rel_err(hz, hz0) < tol
        assert g.keys() == g0.keys()
        for n in g0:
            assert rel_err(g[n], g0[n]) < tol, n


def test_cfg3_training_parity_96_slides():
    """BRCA-shaped survival training step (fwd+bwd) on 96 slides in the default tensor-core mode: large enough for the
    slide-shaped regions (96 -> 128-row stride) to overlap several levels' token regions in the shared scratch buffers, with
    the level sets running concurrently -- values and every live gradient against the oracle."""
    cfg = _config()
    sd = _weights(cfg)
    slides = synth.make_brca_batch(range(96), cfg.num_levels, cfg.model_config.patch_embed_dim)
    ybin, cens = _labels(slides)
    model, rt, hz, loss = _run(cfg, sd, slides, 1, grad=True)
    ocfg = po.OracleConfig.from_config(cfg)
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, 1)


def test_wide_model_dimensions(prec=0):
    """Dimensions beyond the shipped configs': patch_embed_dim 1280 (> 1024: the LSTM backward kernel's multi-column-group
    form), trans_dim 192 with 4 heads of 48 (> 128: two float4 column groups per lane in the importance/token backward,
    the generic-head-width attention kernels), importance hidden 160, fwd+bwd against the oracle.  fp32 mode: the pointwise
    kernels under test are shared by all modes, and the split-bf16 error bound is stated for the shipped K extents (this
    config's 2560-wide gate contraction lands at 1.07e-4 of the gradient scale against the 1e-4 bar)."""
    cfg = _config(num_levels=3, top_k_patches=[6, 6], patch_embed_dim=1280, trans_dim=192, importance_mlp_hidden_dim=160,
                  hierarchical_ctx_mlp_hidden_dim=192)
    sd = _weights(cfg)
    slides = synth.make_brca_batch(range(5), 3, cfg.model_config.patch_embed_dim, wrange=(3, 6), hrange=(2, 5))
    ybin, cens = _labels(slides)
    model, rt, hz, loss = _run(cfg, sd, slides, prec)
    check_against_oracle(po.OracleConfig.from_config(cfg), sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec)
