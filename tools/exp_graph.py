"""Experiment: CUDA-graph replay of the resident step (forward + backward) against eager enqueueing; run under gpurun."""
import os, sys, json
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import BRCA_CONFIG
from paths_b200 import synth
from paths_b200.config import Config
from paths_b200.data_utils.slide import load_patch_preprocessed_slide
from paths_b200.engine import PyramidStore

B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
dev = torch.device("cuda", 0)
cfg = Config.from_dict(BRCA_CONFIG)
torch.manual_seed(cfg.seed)
model = cfg.get_model().to(dev); model.train()
rt = model.runtime(); rt.set_precision(os.environ.get("PREC", "bf16x3"))
D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels
store = PyramidStore(Lv, D, dev)
ids, yb, ce = [], [], []
for sid in range(B):
    s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
    sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(), Lv,
                                       grids=s.grids, store=store)
    ids.append(sl.store_index); yb.append(s.survival_bin); ce.append(s.censored)
caps, maxseg = store.capacities(ids, cfg.top_k_patches)
eng = rt.engine_for(B, caps, maxseg, cfg.top_k_patches)
eng.bind_store(store)
sid_t = torch.tensor(ids, dtype=torch.int32, device=dev)
yb = torch.tensor(yb, dtype=torch.int32, device=dev); ce = torch.tensor(ce, dtype=torch.int32, device=dev)


def step():
    eng.forward(sid_t, yb, ce, 1.0 / B, True)
    eng.backward(True)


def timeit(fn, n=20):
    for _ in range(5):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(n):
        fn()
    b.record(); torch.cuda.synchronize()
    return a.elapsed_time(b) / n


eager = timeit(step)
g = torch.cuda.CUDAGraph()
st = torch.cuda.Stream()
st.wait_stream(torch.cuda.current_stream())
with torch.cuda.stream(st):
    step(); step()
torch.cuda.current_stream().wait_stream(st)
torch.cuda.synchronize()
with torch.cuda.graph(g, stream=st):
    step()
graph = timeit(g.replay)


def grads_after(fn):
    fn(); torch.cuda.synchronize()
    return rt.layout.gflat.clone(), eng.buffer(0, "LOSS").clone()


e1, l1 = grads_after(step); e2, l2 = grads_after(step)
g1, m1 = grads_after(g.replay); g2, m2 = grads_after(g.replay)
sc = float(e1.abs().max())
print(json.dumps({"slides": B, "eager_ms": round(eager, 3), "graph_ms": round(graph, 3),
                  "eager_vs_eager": float((e1 - e2).abs().max()) / sc, "graph_vs_graph": float((g1 - g2).abs().max()) / sc,
                  "graph_vs_eager": float((g1 - e1).abs().max()) / sc, "loss": [float(l1), float(l2), float(m1), float(m2)]}))
off = int((g1 - e1).abs().argmax())
names = [(n, o, k) for n, _, o, k in rt.layout.entries]
print([n for n, o, k in names if o <= off < o + k], off)
