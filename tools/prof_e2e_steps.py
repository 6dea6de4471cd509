"""Per-step wall times of the end-to-end loop on the bench workload, with the pieces switched one at a time (run under gpurun):
    python tools/prof_e2e_steps.py [slides]
Prints, per variant, the per-step milliseconds (host wall clock around one loop iteration, which ends in loss.item())."""
import os, sys, time, json
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
from paths_b200.optim import FlatAdamW

B = int(sys.argv[1]) if len(sys.argv) > 1 else 672
dev = torch.device("cuda", 0)
cfg = Config.from_dict(BRCA_CONFIG)
torch.manual_seed(cfg.seed)
model = cfg.get_model().to(dev); model.train()
rt = model.runtime(); rt.set_precision("bf16x3")
opt = FlatAdamW(model)
D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels
store = PyramidStore(Lv, D, dev)
items = []
for sid in range(B):
    s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
    sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(), Lv,
                                       grids=s.grids, store=store)
    items.append(sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl})
batch = collate_fn(items, packed=True)
batch["fts_packed"] = batch["fts_packed"].pin_memory()
batch["fts"] = batch["fts"][:, :1].clone()
resident = {k: v for k, v in batch.items() if k not in ("fts", "fts_packed")}


def loop(steps, src, prefetch=True):
    it = DevicePrefetcher((src for _ in range(steps)), dev) if prefetch else (src for _ in range(steps))
    ts = []
    t0 = time.perf_counter()
    for b in it:
        opt.zero_grad()
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        loss.backward()
        loss.item()
        t1 = time.perf_counter(); ts.append((t1 - t0) * 1e3); t0 = t1
    return ts


def show(name, ts):
    ts = ts[3:]
    print(f"{name:44s} mean {sum(ts) / len(ts):7.3f}  min {min(ts):7.3f}  max {max(ts):7.3f}   " + " ".join(f"{t:.1f}" for t in ts[:16]), flush=True)


for rep in range(2):
    show("host batch, prefetcher, custom op", loop(23, batch))
    rt.use_custom_op = False
    show("host batch, prefetcher, autograd.Function", loop(23, batch))
    rt.use_custom_op = True
    show("resident store (no H2D), no prefetcher", loop(23, resident, prefetch=False))
    show("resident store (no H2D), prefetcher", loop(23, resident))
    os.environ["PATHS_B200_SKIP_POLL"] = "1"
    show("host batch, prefetcher, no overflow poll", loop(23, batch))
    del os.environ["PATHS_B200_SKIP_POLL"]
torch.cuda.synchronize()
a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
eng = rt.last_engine
ids = torch.tensor([it.store_index for it in batch["slide"]], dtype=torch.int32, device=dev)
yb = batch["survival_bin"].to(dev, torch.int32); ce = batch["censored"].to(dev, torch.int32)
a.record()
for _ in range(20):
    eng.forward(ids, yb, ce, 1.0 / B, True); eng.overflow_poll(); eng.backward(True)
b.record(); torch.cuda.synchronize()
print("resident engine loop ms/step", a.elapsed_time(b) / 20)
