"""TEST / BENCH INFRASTRUCTURE -- drives the UNMODIFIED reference (zzbuzzard/PATHS, imported from a checkout, never copied)
on synthetic pyramids: the recipe of SURVEY.md §8c and tests/golden/make_golden.py as a callable.

The reference is pure Python on PyTorch and is not a package, so it cannot travel to the GPU box; wherever a checkout is
reachable -- this container's /root/reference, or a directory named by $PATHS_REFERENCE -- `bench.py --impl reference` and
`cpu_baseline` time it through this module (kind "reference"), and fall back to the restatement in oracle/paths_oracle.py
(kind "port", pinned to the reference by tests/test_oracle_golden.py) otherwise.  Only tests/, bench.py's reference arm and
its cpu_baseline leg may import this module.
"""
import os
import sys
import tempfile
import types

import torch


def reference_dir():
    for cand in (os.environ.get("PATHS_REFERENCE"), "/root/reference"):
        if cand and os.path.isfile(os.path.join(cand, "model", "paths.py")) and os.path.isfile(os.path.join(cand, "utils.py")):
            return cand
    return None


class ReferenceRunner:
    """fwd(+bwd) of `utils.inference_end2end` of the reference on a fixed batch of synthetic slides."""

    def __init__(self, ref_dir, config_dict, slides):
        os.environ.setdefault("WANDB_MODE", "disabled")
        for name in ["tiatoolbox", "tiatoolbox.wsicore", "tiatoolbox.wsicore.wsireader", "tiatoolbox.tools", "tiatoolbox.tools.tissuemask"]:
            sys.modules.setdefault(name, types.ModuleType(name))          # only the raw-slide path needs tiatoolbox (slide.py:9-10)
        sys.modules["tiatoolbox.wsicore.wsireader"].WSIReader = object
        sys.modules["tiatoolbox.tools.tissuemask"].OtsuTissueMasker = object
        ours = {k: sys.modules.pop(k) for k in list(sys.modules) if k in ("config", "utils") or k.split(".")[0] in ("model", "data_utils", "preprocess")}
        sys.path.insert(0, ref_dir)
        try:
            import config as rconfig
            import utils as rutils
            from data_utils import slide as rslide, dataset as rdataset
            from preprocess import loader as rloader
        finally:
            sys.path.remove(ref_dir)
            sys.modules.update(ours)
        cpu = torch.device("cpu")                                     # the CPU path is what is timed, whatever the host has
        for name, mod in list(sys.modules.items()):
            if (name in ("config", "utils") or name.split(".")[0] in ("model", "data_utils", "preprocess")) and hasattr(mod, "device"):
                mod.device = cpu
        self.rutils = rutils
        data = dict(config_dict)
        mc = dict(data["model_config"])
        data["model_config"] = mc
        cfg = rconfig.Config(**{**data, "model_config": rconfig.PATHSProcessorConfig(**mc),
                                "top_k_patches": ([data["top_k_patches"]] * (data["num_levels"] - 1)
                                                  if isinstance(data["top_k_patches"], int) else list(data["top_k_patches"])),
                                "batch_size": [data["batch_size"]] * data["num_levels"] if isinstance(data["batch_size"], int) else data["batch_size"]})
        cfg.model_config.dropout = 0.0
        self.cfg = cfg
        torch.manual_seed(cfg.seed)
        self.model = cfg.get_model()
        self.model.train()
        tmp = tempfile.mkdtemp(prefix="paths_ref_")
        rloader.set_preprocess_dir(tmp)
        items = []
        for s in slides:
            for l, g in enumerate(s.grids):
                torch.save(g.clone(), os.path.join(tmp, f"s{s.slide_id}_{cfg.base_power * 2 ** l:.3f}.pt"))
            sl = rslide.load_patch_preprocessed_slide(f"/x/s{s.slide_id}.svs", cfg.base_power, cfg.model_config.patch_size,
                                                      self.model.procs[0].ctx_dim(), cfg.num_levels)
            items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl})
        self.items = items
        self.collate = rdataset.collate_fn

    def step(self, backward=True):
        cfg = self.cfg
        batch = self.collate([dict(i) for i in self.items])            # collate_fn pops from its inputs
        if backward:
            self.model.zero_grad(set_to_none=True)
            hz, loss = self.rutils.inference_end2end(cfg.num_levels, cfg.top_k_patches, self.model, cfg.base_power, batch, cfg.task)
            loss.backward()
        else:
            with torch.no_grad():
                hz, loss = self.rutils.inference_end2end(cfg.num_levels, cfg.top_k_patches, self.model, cfg.base_power, batch, cfg.task)
        return float(loss)
