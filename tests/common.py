"""Shared helpers for the tests: rebuild the inputs of a golden fixture, tolerance helpers."""
import os

import torch

from paths_b200 import synth
from oracle import paths_oracle as po

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")

MODEL_DIRS = {
    "sample": dict(),
    "brca_paths_0": dict(),
}
# model_config of models/sample and models/brca_paths_0 (identical apart from root_name; config.json:24-41)
BASE_MODEL_CONFIG = dict(patch_embed_dim=1024, trans_dim=128, trans_heads=4, trans_layers=2,
                         importance_mlp_hidden_dim=128, hierarchical_ctx_mlp_hidden_dim=256, patch_size=256,
                         lstm=True, hierarchical_ctx=True, slide_ctx_mode="residual", importance_mode="mul",
                         pos_encoding_mode="2d")


def load_fixture(name):
    return torch.load(os.path.join(GOLDEN, name + ".pt"), weights_only=False)


def fixture_inputs(fx, grid_edit=None):
    mc = dict(BASE_MODEL_CONFIG)
    mc.update(fx["overrides"])
    mc.pop("dropout", None)
    cfg = po.OracleConfig(num_levels=fx["num_levels"], top_k_patches=fx["top_k"], task="survival", nbins=4, **mc)
    sd = synth.synth_state_dict(fx["shapes"], seed=0)
    slides = []
    for (sid, w0, h0) in fx["slide_specs"]:
        s = synth.make_slide(sid, w0, h0, cfg.num_levels, cfg.patch_embed_dim, fx["p_bg"], seed=fx["base_seed"])
        if grid_edit is not None:
            grid_edit(sid, s.grids)
        slides.append(s)
    ybin = torch.tensor([s.survival_bin for s in slides], dtype=torch.long)
    cens = torch.tensor([s.censored for s in slides], dtype=torch.long)
    return cfg, sd, slides, ybin, cens


def edge_grid_edit(sid, grids):
    # must mirror tests/golden/make_golden.py:kill_children_of_all
    if sid == 1:
        grids[1].zero_()
        grids[1][3, 2] = 1.5
        grids[1][0, 0] = -0.5
    if sid == 2:
        grids[1].zero_()


def rel_err(a, b):
    """max |a-b| / max(|b|, tiny): error relative to the tensor's scale."""
    a = a.double(); b = b.double()
    denom = b.abs().max().clamp(min=1e-30)
    return float((a - b).abs().max() / denom)
