"""CPU-side checks: the C-ABI library loads and exports every symbol include/paths_b200.h declares (no compute calls --
there is no GPU here), and the host-side mirror of the reference API behaves like the reference (config schema,
collate_fn padding, PatchBatch asserts, state_dict keys, flat parameter layout)."""
import ctypes
import os
import re

import pytest
import torch

from tests.common import BASE_CONFIG, product_config, load_fixture, fixture_inputs

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _declared_symbols():
    src = open(os.path.join(ROOT, "include", "paths_b200.h")).read()
    src = re.sub(r"/\*.*?\*/", "", src, flags=re.S)
    return sorted(set(re.findall(r"\b(paths_[a-z0-9_]+)\s*\(", src)))


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as ge
    ge.build()
    from paths_b200 import _lib
    lib = ctypes.CDLL(_lib.LIB_PATH)
    decl = _declared_symbols()
    assert len(decl) >= 20
    for sym in decl:
        assert hasattr(lib, sym), f"{sym} is declared in include/paths_b200.h but not exported"
    assert sorted(_lib.SYMBOLS) == decl, "paths_b200/_lib.py:SYMBOLS must list exactly the declared entry points"
    lib.paths_abi_version.restype = ctypes.c_int
    assert lib.paths_abi_version() == 5


def test_no_cpu_fallback():
    """Without a CUDA device every compute entry fails loudly (engine creation is refused)."""
    if torch.cuda.is_available():
        pytest.skip("needs a machine without a GPU")
    from paths_b200.engine import Engine, make_dims
    cfg = product_config()
    with pytest.raises(RuntimeError, match="no CPU fallback|CUDA"):
        Engine(make_dims(cfg.model_config, cfg), 1, [4] * 5, [4] * 5, "cuda")
    model = cfg.get_model()
    with pytest.raises(RuntimeError, match="CUDA"):
        model.runtime()


def test_product_never_imports_oracle():
    for dirpath, _, files in os.walk(os.path.join(ROOT, "paths_b200")):
        for f in files:
            if f.endswith(".py"):
                src = open(os.path.join(dirpath, f)).read()
                assert not re.search(r"^\s*(from|import)\s+oracle\b", src, flags=re.M), f"{f} imports the oracle"


def test_config_schema_and_state_dict_keys(tmp_path):
    import json
    from paths_b200.config import Config
    d = tmp_path / "m"
    d.mkdir()
    data = dict(BASE_CONFIG)
    data["top_k_patches"] = 20                       # int -> list of L-1 (config.py:93-94)
    (d / "config.json").write_text(json.dumps(data))
    cfg = Config.load(str(d), test_mode=True)
    assert cfg.top_k_patches == [20] * 4 and cfg.num_levels == 5
    model = cfg.get_model()
    keys = set(model.state_dict().keys())
    assert len(keys) and sum(p.numel() for p in model.parameters()) == 9_881_881
    for k in ["lstm.forget_gate.0.weight", "lstm.mem_to_out.0.bias", "procs.0.importance_mlp.2.weight",
              "procs.4.global_agg.special_token", "procs.2.global_agg.transformer.encoder.layers.1.self_attn.in_proj_weight",
              "procs.3.global_agg.transformer.decoder.layers.0.multihead_attn.out_proj.bias",
              "procs.1.global_agg.transformer.decoder.norm.weight", "procs.4.classification_layer.bias"]:
        assert k in keys, k
    assert model.procs[0].ctx_dim() == (128, 1280)
    bad = dict(data); bad["model_type"] = "other"
    (d / "config.json").write_text(json.dumps(bad))
    with pytest.raises(NotImplementedError):
        Config.load(str(d), test_mode=True)


def test_flat_layout_covers_live_parameters():
    from paths_b200.engine import ParamLayout, make_dims
    fx = load_fixture("small_full")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = cfg.get_model()
    model.load_state_dict(sd, strict=True)
    before = {k: v.clone() for k, v in model.state_dict().items()}
    lay = ParamLayout(model, make_dims(cfg.model_config, cfg))
    after = model.state_dict()
    assert all(torch.equal(before[k], after[k]) for k in before), "moving parameters into the flat buffer must not change them"
    live = {n for n, _, _, _ in lay.entries}
    dead = {n for n, _ in lay.dead}
    assert live | dead == {n for n, _ in model.named_parameters()} and not (live & dead)
    assert all(".encoder." in n or "multihead_attn.in_proj" in n or "multihead_attn.out_proj.weight" in n for n in dead), dead
    spans = sorted((o, o + k) for _, _, o, k in lay.entries)
    assert all(a2 >= b1 for (_, b1), (a2, _) in zip(spans, spans[1:])) and spans[-1][1] <= lay.total
    assert lay.valid()


def test_collate_and_patch_batch_match_reference_schema():
    from paths_b200.data_utils.dataset import collate_fn
    from paths_b200.data_utils.patch_batch import PatchBatch, from_batch
    items = []
    for n in (3, 5):
        items.append({"fts": torch.randn(n, 8), "locs": torch.arange(2 * n).view(n, 2) * 256, "parent_inds": torch.arange(n),
                      "ctx_patch": torch.zeros(n, 0, 10), "ctx_slide": torch.zeros(0, 4), "survival_bin": 1, "censored": 0})
    b = collate_fn(items)
    assert b["fts"].shape == (2, 5, 8) and b["ctx_patch"].shape == (2, 5, 0, 10) and b["num_ims"].tolist() == [3, 5]
    assert float(b["fts"][0, 3:].abs().max()) == 0.0
    data = from_batch({k: v for k, v in b.items() if k in ("fts", "locs", "parent_inds", "ctx_patch", "ctx_slide", "num_ims")}, "cpu")
    assert data.valid_inds.sum().item() == 8 and data.ctx_depth == 0 and data.max_patches == 5
    with pytest.raises(AssertionError):              # patch_batch.py:50: padded width must equal the longest bag
        PatchBatch(b["locs"], torch.tensor([3, 4]), b["parent_inds"], b["ctx_slide"], b["ctx_patch"], b["fts"])
