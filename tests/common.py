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
    subtype = fx.get("task") == "subtype_classification"       # nbins carries the number of logits (model/paths.py:25)
    cfg = po.OracleConfig(num_levels=fx["num_levels"], top_k_patches=fx["top_k"], task=fx.get("task", "survival"),
                          nbins=len(fx["subtypes"]) if subtype else 4, **mc)
    sd = synth.synth_state_dict(fx["shapes"], seed=0)
    slides = []
    for (sid, w0, h0) in fx["slide_specs"]:
        s = synth.make_slide(sid, w0, h0, cfg.num_levels, cfg.patch_embed_dim, fx["p_bg"], seed=fx["base_seed"])
        if grid_edit is not None:
            grid_edit(sid, s.grids)
        slides.append(s)
    ybin = torch.tensor([s.survival_bin for s in slides], dtype=torch.long)
    cens = torch.tensor([s.censored for s in slides], dtype=torch.long)
    if subtype:                                                 # the label slot carries the class index, no censoring
        for s, y in zip(slides, fx["subtype_labels"].tolist()):
            s.subtype = int(y)
        return cfg, sd, slides, fx["subtype_labels"].long(), None
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


# ---- product-side helpers (GPU tests) ---------------------------------------------------------------------------
BASE_CONFIG = {   # models/sample/config.json and models/brca_paths_0/config.json of the reference (identical but root_name)
    "wsi_dir": "/x", "preprocess_dir": "/x", "csv_path": "/x", "seed": 0, "batch_size": 32, "save_epochs": 10,
    "eval_epochs": 1, "lr": 2e-5, "lr_decay_per_epoch": 0.99, "base_power": 0.625, "magnification_factor": 2,
    "num_levels": 5, "num_epochs": 40, "min_epochs": 0, "top_k_patches": 20, "early_stopping": False,
    "hipt_splits": True, "hipt_val_proportion": 0.0, "model_type": "PATHS",
    "model_config": {"hierarchical_ctx": True, "slide_ctx_mode": "residual", "patch_embed_dim": 1024, "dropout": 0.0,
                     "importance_mode": "mul", "trans_dim": 128, "trans_heads": 4, "trans_layers": 2,
                     "pos_encoding_mode": "2d", "importance_mlp_hidden_dim": 128,
                     "hierarchical_ctx_mlp_hidden_dim": 256, "lstm": True},
    "root_name": "sample",
}


def product_config(fx=None, **over):
    import copy
    from paths_b200.config import Config
    data = copy.deepcopy(BASE_CONFIG)
    if fx is not None:
        data["model_config"].update(fx["overrides"])
        data["num_levels"] = fx["num_levels"]
        data["top_k_patches"] = list(fx["top_k"])
        if fx.get("task") == "subtype_classification":
            data["task"] = fx["task"]
            data["filter_to_subtypes"] = list(fx["subtypes"])
    data["model_config"]["dropout"] = 0.0
    for k, v in over.items():
        if k in data["model_config"]:
            data["model_config"][k] = v
        else:
            data[k] = v
    return Config.from_dict(data, test_mode=True)


def product_model(cfg, sd, device="cuda"):
    model = cfg.get_model()
    missing = model.load_state_dict(sd, strict=True)
    return model.to(device)


def product_batch(cfg, model, slides, store=None, device="cuda"):
    from paths_b200.data_utils.slide import load_patch_preprocessed_slide
    from paths_b200.data_utils.dataset import collate_fn
    from paths_b200.engine import PyramidStore
    if store is None:
        store = PyramidStore(cfg.num_levels, cfg.model_config.patch_embed_dim, device)
    items = []
    for s in slides:
        sl = load_patch_preprocessed_slide(f"/x/s{s.slide_id}.svs", cfg.base_power, cfg.model_config.patch_size,
                                           model.procs[0].ctx_dim(), cfg.num_levels, grids=s.grids, store=store,
                                           subtype=getattr(s, "subtype", None))           # todict() adds it (slide.py:367-368)
        items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored,
                                    "slide": sl})
    return collate_fn(items), store
