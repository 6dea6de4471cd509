"""CPU-only companion of tools/parity_report.py: the float64 ORACLE against itself with every weight perturbed by a relative\n2^-17 (the per-product rounding of the three-term bf16 split), brca_b4 fixture, teacher-forced selection.  Shows how far the\nmodel s own gradients move, tensor by tensor, under a perturbation of that size (ReLU kinks flip).  python tools/parity_perturbation.py"""
import sys; import os; sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import torch
from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, rel_err
fx = load_fixture("brca_b4")
ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
grids=[[g.double() for g in s.grids] for s in slides]
sd64={k: v.double() if v.is_floating_point() else v for k,v in sd.items()}
_,_,recs = po.end2end(sd64, ocfg, grids, ybin, cens, dtype=torch.float64, record=True)
forced=[[k for k in recs[l].keep_inds] for l in range(ocfg.num_levels-1)]
_,_,g64,_ = po.grads(sd64, ocfg, grids, ybin, cens, dtype=torch.float64, forced_keep=forced)
for seed in (1,2,3):
    gp = torch.Generator().manual_seed(seed)
    sdp = {k: (v * (1 + 2.0 ** -17 * (2 * torch.rand(v.shape, generator=gp, dtype=torch.float64) - 1)) if v.is_floating_point() else v) for k, v in sd64.items()}
    _,_,gp_,_ = po.grads(sdp, ocfg, grids, ybin, cens, dtype=torch.float64, forced_keep=forced)
    errs = {k: rel_err(gp_[k], g64[k]) for k in g64 if g64[k] is not None and float(g64[k].abs().max())>0}
    top = sorted(errs.items(), key=lambda kv:-kv[1])[:6]
    print(seed, "n>1e-4:", sum(e>1e-4 for e in errs.values()), [(k[-40:], f"{e:.1e}") for k,e in top])
