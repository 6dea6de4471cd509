"""Per-tensor parity report of the CUDA path against the teacher-forced CPU oracle (run on the B200 box).
Prints, for every precision mode, the worst value errors per level and the gradient errors of every parameter tensor in
two norms: max|a-b| / max|b| and ||a-b||_2 / ||b||_2."""
import sys, os, json
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, product_config, product_model, product_batch, rel_err
from paths_b200.utils import inference_end2end

name = sys.argv[1] if len(sys.argv) > 1 else "brca_b4"
out = {}
for prec, pname in [(0, "fp32"), (1, "bf16x3"), (2, "bf16")]:
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    rt = model.runtime(); rt.set_precision(prec); rt.strict_overflow = True
    batch, store = product_batch(cfg, model, slides)
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward(); torch.cuda.synchronize()
    eng = rt.last_engine; B = len(slides)
    forced = []
    for l in range(cfg.num_levels - 1):
        seg = eng.buffer(l, "SEG")[:B + 1].tolist()
        kcap = eng.buffer(l, "SEL").numel() // B
        sel = eng.buffer(l, "SEL").view(B, kcap).cpu(); cnt = eng.buffer(l, "SEL_COUNT")[:B].cpu()
        forced.append([(sel[j, :int(cnt[j])] - seg[j]).long() for j in range(B)])
    ohz, oloss, og, recs = po.grads(sd, ocfg, [s.grids for s in slides], ybin, cens, forced_keep=forced)
    rep = {"hazards": rel_err(hz.detach().cpu(), ohz), "loss": abs(float(loss) - float(oloss)) / abs(float(oloss))}
    for l in range(cfg.num_levels):
        M = int(eng.buffer(l, "COUNT").item())
        a = eng.buffer(l, "ALPHA")[:M].cpu()
        ref = torch.cat([recs[l].importance[j, :int(recs[l].num_ims[j])] for j in range(B)])
        rep[f"importance_l{l}"] = rel_err(a, ref)
        rep[f"logits_l{l}"] = rel_err(eng.buffer(l, "LOGITS").view(B, -1).cpu(), recs[l].logits)
    named = dict(model.named_parameters())
    gm, gl = {}, {}
    for k, g in og.items():
        if g is None or float(g.abs().max()) == 0: continue
        mine = named[k].grad.detach().cpu().double(); g = g.double()
        gm[k] = float((mine - g).abs().max() / g.abs().max()); gl[k] = float((mine - g).norm() / g.norm())
    rep["grad_max_norm_worst"] = sorted(gm.items(), key=lambda kv: -kv[1])[:6]
    rep["grad_l2_worst"] = sorted(gl.items(), key=lambda kv: -kv[1])[:6]
    out[pname] = rep
    print(pname, json.dumps(rep, indent=1))
