"""Where the end-to-end step loop spends host time between the loss read-back of one step and the first launch of the next
(the GPU idles there); run under gpurun:  python tools/prof_e2e_gap.py [slides]"""
import os, sys, time
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
import paths_b200.engine as E

B = int(sys.argv[1]) if len(sys.argv) > 1 else 672
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

marks = {}
orig_fwd = E.Engine.forward
def fwd(self, *a, **k):
    marks["fwd_in"] = time.perf_counter()
    r = orig_fwd(self, *a, **k)
    marks["fwd_out"] = time.perf_counter()
    return r
E.Engine.forward = fwd
orig_bwd = E.Engine.backward
def bwd(self, *a, **k):
    marks["bwd_in"] = time.perf_counter()
    r = orig_bwd(self, *a, **k)
    marks["bwd_out"] = time.perf_counter()
    return r
E.Engine.backward = bwd

acc = {}
def add(k, v): acc[k] = acc.get(k, 0.0) + v

def run(steps, rec):
    t_item = None
    it = iter(DevicePrefetcher((batch for _ in range(steps)), dev))
    while True:
        t0 = time.perf_counter()
        try:
            b = next(it)
        except StopIteration:
            break
        t1 = time.perf_counter()
        opt.zero_grad()
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        t2 = time.perf_counter()
        loss.backward()
        t3 = time.perf_counter()
        loss.item()
        t4 = time.perf_counter()
        if rec:
            add("prefetch_next", t1 - t0); add("e2e_python_before_engine", marks["fwd_in"] - t1); add("engine_forward_enqueue", marks["fwd_out"] - marks["fwd_in"])
            add("after_forward_python", t2 - marks["fwd_out"]); add("autograd_before_engine", marks["bwd_in"] - t2)
            add("engine_backward_enqueue", marks["bwd_out"] - marks["bwd_in"]); add("after_backward_python", t3 - marks["bwd_out"])
            add("item_wait", t4 - t3); add("step", t4 - t0)

run(5, False)
torch.cuda.synchronize()
N = 20
run(N, True)
for k, v in acc.items():
    print(f"{k:32s} {v / N * 1e3:8.3f} ms")
