"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
fp32 mode tolerance: 1e-4 relative to each tensor's scale (BASELINE.json north_star); indices bit-exact wherever the
oracle's scores are separated by more than the tolerance; values compared with teacher-forced indices."""
import os

import pytest
import torch

from oracle import paths_oracle as po
from tests.common import (load_fixture, fixture_inputs, edge_grid_edit, rel_err, product_config, product_model,
                          product_batch)

from tests.parity_check import check_against_oracle, TOL

pytestmark = pytest.mark.gpu
PREC_NAMES = {0: "fp32", 1: "bf16x3", 2: "bf16"}


def _gemm(prec, A, B, b_kn, bias=None, flags=0, C0=None):
    from paths_b200 import _lib as L
    lib = L.lib()
    M, K = A.shape
    N = B.shape[1] if b_kn else B.shape[0]
    C = torch.zeros(M, N, device="cuda") if C0 is None else C0.clone()
    L.check(lib.paths_gemm_nt(prec, A.data_ptr(), A.stride(0), B.data_ptr(), B.stride(0), int(b_kn), C.data_ptr(), C.stride(0),
                              bias.data_ptr() if bias is not None else None, M, N, K, flags,
                              torch.cuda.current_stream().cuda_stream), "gemm_nt")
    return C


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("shape", [(1, 4, 4), (77, 136, 88), (300, 1792, 2048), (513, 256, 1024), (129, 128, 512), (1000, 384, 128)])
def test_gemm_nt(prec, shape):
    M, N, K = shape
    g = torch.Generator(device="cuda").manual_seed(M * 7 + N)
    A = torch.randn(M, K, device="cuda", generator=g)
    W = torch.randn(N, K, device="cuda", generator=g) / K ** 0.5
    bias = torch.randn(N, device="cuda", generator=g)
    ref = (A.double() @ W.double().t() + bias.double())
    out = _gemm(prec, A, W, False, bias)
    assert rel_err(out, ref) < TOL[prec]
    out = _gemm(prec, A, W.t().contiguous(), True, bias, flags=1)
    assert rel_err(out, torch.relu(ref)) < TOL[prec]
    C0 = torch.randn(M, N, device="cuda", generator=g)
    out = _gemm(prec, A, W, False, None, flags=2, C0=C0)
    assert rel_err(out, ref - bias.double() + C0.double()) < TOL[prec]


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("shape", [(5, 4, 8), (777, 136, 88), (4000, 1792, 1024), (3001, 4, 128), (2500, 512, 128)])
def test_gemm_tn(prec, shape):
    from paths_b200 import _lib as L
    lib = L.lib()
    M, NI, NJ = shape
    g = torch.Generator(device="cuda").manual_seed(M + NI)
    P = torch.randn(M, NI, device="cuda", generator=g)
    Q = torch.randn(M, NJ, device="cuda", generator=g)
    C = torch.zeros(NI, NJ, device="cuda")
    L.check(lib.paths_gemm_tn(prec, P.data_ptr(), NI, Q.data_ptr(), NJ, C.data_ptr(), NJ, M, NI, NJ,
                              torch.cuda.current_stream().cuda_stream), "gemm_tn")
    ref = P.double().t() @ Q.double()
    assert rel_err(C, ref) < TOL[prec]


def _run_product(name, prec):
    fx = load_fixture(name)
    edit = edge_grid_edit if name == "small_edge" else None
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx, edit)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    rt = model.runtime()
    rt.set_precision(prec)
    batch, store = product_batch(cfg, model, slides)
    from paths_b200.utils import inference_end2end
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    torch.cuda.synchronize()
    return fx, ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("name", ["small_full", "small_edge", "cfg1_sample", "brca_b4", "small_rnn", "small_concat1d", "small_subtype"])
def test_end2end_parity(name, prec):
    fx, ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss = _run_product(name, prec)
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec)
    if name == "small_subtype":                      # task "subtype_classification": logits + cross-entropy (utils.py:274-278)
        assert torch.allclose(hz.detach().cpu(), fx["hazards"], atol=TOL[prec] * float(fx["hazards"].abs().max()))
        assert abs(float(loss) - float(fx["loss"])) < TOL[prec] * 2


def test_risk_scores_match_the_evaluator():
    """eval.py:60-61: risk = -sum(cumprod(1 - hazards)); the loss kernel writes it next to the hazards."""
    fx, ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss = _run_product("brca_b4", 0)
    want = -torch.sum(torch.cumprod(1 - hz.detach(), dim=1), dim=1)
    assert rt.risk().shape == (len(slides),) and torch.allclose(rt.risk(), want, rtol=1e-6, atol=1e-6)
    # ... and against the oracle's hazards (teacher-forced with our selection: level 0 holds exact importance ties among its
    # background cells, whose top-K order torch leaves unspecified)
    eng = rt.last_engine
    B = len(slides)
    forced = []
    for l in range(cfg.num_levels - 1):
        seg = eng.buffer(l, "SEG")[:B + 1].tolist()
        kcap = eng.buffer(l, "SEL").numel() // B
        sel = eng.buffer(l, "SEL").view(B, kcap).cpu(); cnt = eng.buffer(l, "SEL_COUNT")[:B].cpu()
        forced.append([(sel[j, :int(cnt[j])] - seg[j]).long() for j in range(B)])
    ohz, _, _ = po.end2end(sd, ocfg, [s.grids for s in slides], ybin, cens, forced_keep=forced)
    assert rel_err(rt.risk().cpu(), -torch.sum(torch.cumprod(1 - ohz, dim=1), dim=1)) < 1e-4


def fallback_in_slack_batch(fx, ocfg, sd, n_slides=24, victim=7):
    """24 two-by-two slides (small_edge model: K = [8, 2, -1]).  Level 2 of every slide is edited after a first oracle pass
    has told which two level-1 rows each slide keeps: ordinary slides keep ONE tissue cell under their best parent (1 row of
    their 8-row capacity: the batch-wide slack), the victim loses every cell under both kept parents and therefore takes the
    no-child fallback with the 56 remaining cells -- above the per-slide bound 8, inside the batch total."""
    import paths_b200.synth as synth
    slides = [synth.make_slide(100 + i, 2, 2, ocfg.num_levels, ocfg.patch_embed_dim, 0.0, seed=fx["base_seed"]) for i in range(n_slides)]
    ybin = torch.tensor([s.survival_bin for s in slides]); cens = torch.tensor([s.censored for s in slides])
    _, _, free = po.end2end(sd, ocfg, [s.grids for s in slides], ybin, cens, record=True)
    for j, s in enumerate(slides):
        keep = free[1].keep_inds[j]
        xy = (free[1].locs[j][keep] // ocfg.patch_size).tolist()          # kept level-1 cells, importance-descending
        g2 = s.grids[2]
        if j == victim:
            for x, y in xy:
                g2[2 * x:2 * x + 2, 2 * y:2 * y + 2] = 0.0
        else:
            one = g2[2 * xy[0][0], 2 * xy[0][1]].clone()
            g2.zero_()
            g2[2 * xy[0][0], 2 * xy[0][1]] = one
    return slides, ybin, cens


def test_fallback_slide_inside_a_large_batch():
    """A slide that takes the no-child fallback (slide.py:336-352) inside a batch whose capacity slack would hide it: the bag
    exceeds the level's per-slide bound although the batch total fits.  The step must be re-run on worst-case buffers before
    any output or gradient exists, and match the oracle."""
    fx = load_fixture("small_edge")
    ocfg, sd, _, _, _ = fixture_inputs(fx, edge_grid_edit)
    slides, ybin, cens = fallback_in_slack_batch(fx, ocfg, sd)
    _, _, free = po.end2end(sd, ocfg, [s.grids for s in slides], ybin, cens, record=True)
    assert free[1].fallback[7] and sum(free[1].fallback) == 1
    bags = free[2].num_ims.tolist()
    assert bags[7] == 56 and sum(bags) <= 8 * len(slides), "the fallback bag must fit the batch total but not the per-slide bound"
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    rt = model.runtime(); rt.set_precision(0)
    batch, store = product_batch(cfg, model, slides)
    from paths_b200.utils import inference_end2end
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    torch.cuda.synchronize()
    assert rt.overflow_reruns == 1
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, 0)


def test_no_gradient_from_an_overflowed_pass():
    """The truncated pass of an overflowing step never reaches autograd: with the worst-case re-run made impossible the call
    raises before returning, and no parameter has a .grad."""
    fx = load_fixture("small_edge")
    ocfg, sd, _, _, _ = fixture_inputs(fx, edge_grid_edit)
    slides, ybin, cens = fallback_in_slack_batch(fx, ocfg, sd)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    rt = model.runtime(); rt.set_precision(0)
    batch, store = product_batch(cfg, model, slides)
    real = store.capacities
    store.capacities = lambda ids, top_k, worst_case=False: real(ids, top_k, worst_case=False)    # never grant the worst case
    from paths_b200.utils import inference_end2end
    with pytest.raises(RuntimeError, match="overflow"):
        inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    assert all(p.grad is None for p in model.parameters())


@pytest.mark.parametrize("name", ["small_full", "cfg1_sample", "small_rnn", "small_concat1d"])
def test_level_api_matches_oracle(name):
    """RecursiveModel.forward(depth, PatchBatch) on the oracle's own padded batches (reference API, padded layout)."""
    from paths_b200.data_utils.patch_batch import PatchBatch
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    batch = po.collate([po.level0_items(s.grids, ocfg) for s in slides])
    for depth in range(ocfg.num_levels):
        want = po.level_forward(sd, ocfg, depth, batch)
        data = PatchBatch(**{k: v.cuda() for k, v in batch.items()})
        got = model(depth, data)
        for key in ("logits", "ctx_slide", "importance", "ctx_patch"):
            assert rel_err(got[key].cpu(), want[key] * (1 if key != "ctx_patch" else
                           (torch.arange(batch["fts"].shape[1])[None] < batch["num_ims"][:, None])[..., None])) < 1e-4, (depth, key)
        if depth == ocfg.num_levels - 1:
            break
        items = []
        for j, s in enumerate(slides):
            n = int(batch["num_ims"][j])
            items.append(po.select_expand(s.grids[depth + 1], n, batch["locs"][j], batch["ctx_slide"][j], batch["ctx_patch"][j],
                                          want["importance"][j], want["ctx_slide"][j], want["ctx_patch"][j],
                                          ocfg.top_k_patches[depth], ocfg.patch_size))
        batch = po.collate(items)
