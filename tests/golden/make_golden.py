"""Generates tests/golden/*.pt by running the UNMODIFIED reference (imported from /root/reference; never
copied) on seeded synthetic pyramids with deterministic weights.  Run in the build container only:

    python tests/golden/make_golden.py

The GPU box has no /root/reference; tests read the committed fixtures.  Import recipe: SURVEY.md §8(c)
(tiatoolbox is absent and only needed by the raw-slide path, so five empty modules stand in for it).
"""
import os
import sys
import types
import tempfile

os.environ["WANDB_MODE"] = "disabled"
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
REF = os.environ.get("PATHS_REFERENCE", "/root/reference")
for name in ["tiatoolbox", "tiatoolbox.wsicore", "tiatoolbox.wsicore.wsireader", "tiatoolbox.tools",
             "tiatoolbox.tools.tissuemask"]:
    sys.modules[name] = types.ModuleType(name)
sys.modules["tiatoolbox.wsicore.wsireader"].WSIReader = object
sys.modules["tiatoolbox.tools.tissuemask"].OtsuTissueMasker = object
sys.path.insert(0, REF)

import torch  # noqa: E402
import config as rconfig  # noqa: E402
import utils as rutils  # noqa: E402
from data_utils import slide as rslide, dataset as rdataset  # noqa: E402
from preprocess import loader as rloader  # noqa: E402

sys.path.insert(1, ROOT)
from paths_b200 import synth  # noqa: E402


def sample_idx(numel, n=48, seed=0):
    g = torch.Generator().manual_seed(seed)
    return torch.randint(0, numel, (min(n, numel),), generator=g)


def run_case(name, model_dir, slide_specs, overrides=None, top_k=None, num_levels=None, p_bg=0.3,
             grid_edit=None, full_grads=False, base_seed=1000, subtypes=None):
    """slide_specs: list of (slide_id, w0, h0).  subtypes: class names => task "subtype_classification" (utils.py:274-278),
    slide i labelled i mod len(subtypes) (dataset.py:161-163 stores the index into config.filter_to_subtypes)."""
    if only and name not in only:
        return
    cfg = rconfig.Config.load(os.path.join(REF, "models", model_dir), test_mode=True)
    cfg.model_config.dropout = 0.0
    if subtypes is not None:
        cfg.task = "subtype_classification"
        cfg.filter_to_subtypes = list(subtypes)
    for k, v in (overrides or {}).items():
        setattr(cfg.model_config, k, v)
    if num_levels is not None:
        cfg.num_levels = num_levels
    if top_k is not None:
        cfg.top_k_patches = list(top_k)
    else:
        cfg.top_k_patches = cfg.top_k_patches[:cfg.num_levels - 1]
    torch.manual_seed(cfg.seed)
    model = cfg.get_model()
    shapes = {k: tuple(v.shape) for k, v in model.state_dict().items()}
    model.load_state_dict(synth.synth_state_dict(shapes, seed=0))
    model.train()   # dropout is 0, so train == eval numerically; grads flow

    D = cfg.model_config.patch_embed_dim
    tmp = tempfile.mkdtemp()
    rloader.set_preprocess_dir(tmp)
    items = []
    for (sid, w0, h0) in slide_specs:
        s = synth.make_slide(sid, w0, h0, cfg.num_levels, D, p_bg, seed=base_seed)
        if grid_edit is not None:
            grid_edit(sid, s.grids)
        for l, g in enumerate(s.grids):
            torch.save(g.clone(), os.path.join(tmp, f"s{sid}_{cfg.base_power * 2 ** l:.3f}.pt"))
        sl = rslide.load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size,
                                                  model.procs[0].ctx_dim(), cfg.num_levels,
                                                  subtype=None if subtypes is None else sid % len(subtypes))
        items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival,
                                    "censored": s.censored, "slide": sl})
    batch = rdataset.collate_fn(items)

    levels = []
    orig_forward = model.forward

    def rec_forward(depth, data, **kw):
        out = orig_forward(depth, data, **kw)
        levels.append({"num_ims": data.num_ims.clone(), "locs": data.locs.clone(),
                       "parent_inds": data.parent_inds.clone(),
                       "logits": out["logits"].detach().clone(), "ctx_slide": out["ctx_slide"].detach().clone(),
                       "importance": out["importance"].detach().clone(),
                       "ctx_patch_head": out["ctx_patch"][:, :4].detach().clone(),
                       "ctx_patch_norm": out["ctx_patch"].detach().double().norm().item()})
        return out
    model.forward = rec_forward
    hz, loss = rutils.inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    grads = {}
    for k, p in model.named_parameters():
        if p.grad is None:
            grads[k] = None
        else:
            g = p.grad.detach()
            idx = sample_idx(g.numel(), seed=len(k))
            grads[k] = {"norm": g.double().norm().item(), "sum": g.double().sum().item(),
                        "idx": idx, "vals": g.reshape(-1)[idx].clone()}
            if full_grads:
                grads[k]["full"] = g.clone()
    fixture = {
        "name": name, "model_dir": model_dir, "overrides": overrides or {}, "top_k": list(cfg.top_k_patches),
        "num_levels": cfg.num_levels, "slide_specs": slide_specs, "p_bg": p_bg, "base_seed": base_seed,
        "shapes": shapes, "levels": levels, "hazards": hz.detach().clone(), "loss": loss.detach().clone(),
        "grads": grads, "torch": torch.__version__,
    }
    if subtypes is not None:
        fixture["task"] = cfg.task
        fixture["subtypes"] = list(subtypes)
        fixture["subtype_labels"] = batch["subtype"].clone()
    torch.save(fixture, os.path.join(HERE, f"{name}.pt"))
    print(name, "loss", float(loss), "levels", [int(l["num_ims"].sum()) for l in levels])


def kill_children_of_all(sid, grids):
    """Edge case slide.py:336-352: every level-1 cell is background for slide 1 -> fallback path, and
    level 2 is all background too (the `filter[:] = True` branch) for slide 2."""
    if sid == 1:
        grids[1].zero_()
        grids[1][3, 2] = 1.5
        grids[1][0, 0] = -0.5
    if sid == 2:
        grids[1].zero_()


SMALL = {"patch_embed_dim": 64, "trans_dim": 32, "importance_mlp_hidden_dim": 16,
         "hierarchical_ctx_mlp_hidden_dim": 24}

only = set(sys.argv[1:])      # python make_golden.py [case ...]: regenerate only the named fixtures

if __name__ == "__main__":
    # cfg 1 of BASELINE.json: models/sample, one 8x7 slide, L=5, K=20
    run_case("cfg1_sample", "sample", [(0, 8, 7)])
    # 4 BRCA-shaped slides, brca_paths_0 config, fwd+bwd (ragged bags, padding)
    run_case("brca_b4", "brca_paths_0", [(i, *synth.brca_shape(i)) for i in range(4)])
    # small-dim model with full gradients committed
    run_case("small_full", "sample", [(i, *synth.brca_shape(i, wrange=(2, 5), hrange=(2, 4))) for i in range(3)],
             overrides=SMALL, top_k=[3, 5, -1], num_levels=4, full_grads=True)
    # edge cases: fallback path, K > n, K = -1
    run_case("small_edge", "sample", [(0, 2, 2), (1, 2, 3), (2, 1, 2)], overrides=SMALL, top_k=[8, 2, -1],
             num_levels=4, grid_edit=kill_children_of_all, full_grads=True)
    # variants (oracle-only coverage)
    run_case("small_rnn", "sample", [(0, 3, 2), (1, 2, 2)], overrides=SMALL | {"lstm": False}, top_k=[3, 3], num_levels=3,
             full_grads=True)
    run_case("small_concat1d", "sample", [(0, 3, 2), (1, 2, 2)],
             overrides=SMALL | {"slide_ctx_mode": "concat", "pos_encoding_mode": "1d", "importance_mode": "none"},
             top_k=[3, 3], num_levels=3, full_grads=True)
    # task "subtype_classification" (utils.py:274-278: cross-entropy on the last level's logits, three classes)
    run_case("small_subtype", "sample", [(i, *synth.brca_shape(i, wrange=(2, 4), hrange=(2, 4))) for i in range(4)],
             overrides=SMALL, top_k=[4, 4], num_levels=3, full_grads=True, subtypes=["IDC", "ILC", "MDLC"])
