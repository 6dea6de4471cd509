"""cProfile of the end-to-end step loop (host side) on the bench workload; run under gpurun."""
import cProfile, pstats, os, sys, io
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import BRCA_CONFIG
from paths_b200 import synth
from paths_b200.config import Config
from paths_b200.data_utils.dataset import collate_fn
from paths_b200.data_utils.prefetch import DevicePrefetcher
from paths_b200.data_utils.slide import load_patch_preprocessed_slide
from paths_b200.engine import PyramidStore
from paths_b200.utils import inference_end2end

B = int(sys.argv[1]) if len(sys.argv) > 1 else 256
dev = torch.device("cuda", 0)
cfg = Config.from_dict(BRCA_CONFIG)
torch.manual_seed(cfg.seed)
model = cfg.get_model().to(dev); model.train()
rt = model.runtime(); rt.set_precision("bf16x3")
from paths_b200.optim import FlatAdamW
opt = FlatAdamW(model)
D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels
store = PyramidStore(Lv, D, dev)
items = []
for sid in range(B):
    s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
    sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(), Lv,
                                       grids=s.grids, store=store)
    items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl})
batch = collate_fn(items)
batch["fts"] = batch["fts"].pin_memory()


def run(steps):
    for b in DevicePrefetcher((batch for _ in range(steps)), dev):
        opt.zero_grad()
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        loss.backward()
        loss.item()


run(5)
torch.cuda.synchronize()
import time
t0 = time.perf_counter(); run(20); torch.cuda.synchronize(); print("ms/step", (time.perf_counter() - t0) * 50)
pr = cProfile.Profile(); pr.enable(); run(20); pr.disable()
s = io.StringIO(); pstats.Stats(pr, stream=s).sort_stats("cumulative").print_stats(28); print(s.getvalue()[:6000])
