"""Experiment: K independent micro-batch engines on separate caller streams (pointwise kernels of one overlap the contractions
of another).  Timing only -- the engines share the packed-weight registry, which is only safe while the weights are frozen."""
import os, sys, time, json, ctypes as C
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import torch
from bench import BRCA_CONFIG
from paths_b200 import synth, _lib as L
from paths_b200.config import Config
from paths_b200.data_utils.slide import load_patch_preprocessed_slide
from paths_b200.engine import PyramidStore


def main():
    Btot = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    K = int(sys.argv[2]) if len(sys.argv) > 2 else 2
    steps, warm = 20, 5
    dev = torch.device("cuda", 0)
    cfg = Config.from_dict(BRCA_CONFIG)
    torch.manual_seed(cfg.seed)
    model = cfg.get_model().to(dev); model.train()
    rt = model.runtime(); rt.set_precision(os.environ.get("PREC", "bf16x3"))
    D, Lv = cfg.model_config.patch_embed_dim, cfg.num_levels
    store = PyramidStore(Lv, D, dev)
    slides = []
    for sid in range(Btot):
        s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
        sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(), Lv,
                                           grids=s.grids, store=store)
        slides.append((sl, s))
    per = Btot // K
    engs = []
    for k in range(K):
        part = slides[k * per:(k + 1) * per]
        ids = [sl.store_index for sl, _ in part]
        caps, maxseg = store.capacities(ids, cfg.top_k_patches)
        eng = rt.engine_for(per, caps, maxseg, cfg.top_k_patches, tag=f"mb{k}")
        eng.bind_store(store)
        g = rt.layout.gflat if k == 0 else torch.zeros_like(rt.layout.gflat)
        L.check(eng.lib.paths_engine_bind_params(eng.handle, rt.layout.flat.data_ptr(), g.data_ptr(), rt.layout.total, rt.layout.offsets,
                                                 rt.div.data_ptr()), "bind")
        sid_t = torch.tensor(ids, dtype=torch.int32, device=dev)
        yb = torch.tensor([s.survival_bin for _, s in part], dtype=torch.int32, device=dev)
        ce = torch.tensor([s.censored for _, s in part], dtype=torch.int32, device=dev)
        engs.append((eng, sid_t, yb, ce, g, torch.cuda.Stream(dev)))
    scale = 1.0 / Btot

    def step():
        cur = torch.cuda.current_stream()
        evs = []
        for eng, sid_t, yb, ce, g, st in engs:
            st.wait_stream(cur)
            with torch.cuda.stream(st):
                eng.forward(sid_t, yb, ce, scale, True)
        for eng, sid_t, yb, ce, g, st in engs:
            with torch.cuda.stream(st):
                eng.backward(True)
        for eng, sid_t, yb, ce, g, st in engs:
            cur.wait_stream(st)
        for k in range(1, K):
            engs[0][4].add_(engs[k][4])

    for _ in range(warm):
        step()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(steps):
        step()
    b.record(); torch.cuda.synchronize()
    ms = a.elapsed_time(b) / steps
    print(json.dumps({"slides": Btot, "microbatches": K, "ms_per_step": round(ms, 3), "slides_per_s": round(Btot / ms * 1e3)}))


if __name__ == "__main__":
    main()
