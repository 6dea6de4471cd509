"""GPU parity tests: the CUDA path (through the C ABI) against the CPU oracle on the same seeded inputs.
fp32 mode tolerance: 1e-4 relative to each tensor's scale (BASELINE.json north_star); indices bit-exact wherever the
oracle's scores are separated by more than the tolerance; values compared with teacher-forced indices."""
import os

import pytest
import torch

from oracle import paths_oracle as po
from tests.common import (load_fixture, fixture_inputs, edge_grid_edit, rel_err, product_config, product_model,
                          product_batch)

pytestmark = pytest.mark.gpu
TOL = {0: 1e-4, 1: 1e-4, 2: 2e-2}
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
    rt.strict_overflow = True
    batch, store = product_batch(cfg, model, slides)
    from paths_b200.utils import inference_end2end
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    torch.cuda.synchronize()
    return fx, ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss


def _level_rows(eng, l, B):
    M = int(eng.buffer(l, "COUNT").item())
    seg = eng.buffer(l, "SEG")[:B + 1].tolist()
    return M, seg


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("name", ["small_full", "small_edge", "cfg1_sample", "brca_b4", "small_rnn", "small_concat1d"])
def test_end2end_parity(name, prec):
    fx, ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss = _run_product(name, prec)
    eng = rt.last_engine
    tol = TOL[prec]
    B, L = len(slides), cfg.num_levels
    assert not eng.overflowed()[0]

    # 1. free-running oracle: our selection must be a valid top-K of the oracle's scores
    _, _, free = po.end2end(sd, ocfg, [s.grids for s in slides], ybin, cens, record=True)
    forced = []
    for l in range(L - 1):
        M, seg = _level_rows(eng, l, B)
        kcap = eng.buffer(l, "SEL").numel() // B
        sel = eng.buffer(l, "SEL").view(B, kcap).cpu()
        cnt = eng.buffer(l, "SEL_COUNT")[:B].cpu()
        forced.append([(sel[j, :int(cnt[j])] - seg[j]).long() for j in range(B)])
    # 2. teacher-forced oracle (our indices) for value comparisons
    ohz, oloss, ograds, recs = po.grads(sd, ocfg, [s.grids for s in slides], ybin, cens, forced_keep=forced)
    for l in range(L):
        M, seg = _level_rows(eng, l, B)
        rec = recs[l]
        assert seg[1:] == torch.cumsum(rec.num_ims, 0).tolist(), f"level {l}: bag sizes differ"
        xy = eng.buffer(l, "ROW_XY").view(-1, 2)[:M].cpu().long()
        alpha = eng.buffer(l, "ALPHA")[:M].cpu()
        H = eng.buffer(l, "H").view(-1, ocfg.patch_embed_dim)[:M].cpu()
        Cn = eng.buffer(l, "CN").view(-1, ocfg.hierarchical_ctx_mlp_hidden_dim)[:M].cpu()
        for j in range(B):
            n = int(rec.num_ims[j])
            sl = slice(seg[j], seg[j] + n)
            assert torch.equal(xy[sl] * ocfg.patch_size, rec.locs[j, :n]), f"level {l} slide {j}: child locations differ"
            assert rel_err(alpha[sl], rec.importance[j, :n]) < tol, f"level {l} slide {j}: importance"
            state = torch.cat((H[sl], Cn[sl]), 1) if ocfg.lstm else H[sl]          # RNN variant: patch state = Z
            assert rel_err(state, rec.ctx_patch[j, :n]) < tol, f"level {l} slide {j}: patch state"
        assert rel_err(eng.buffer(l, "LOGITS").view(B, -1).cpu(), rec.logits) < tol, f"level {l}: logits"
        assert rel_err(eng.buffer(l, "F").view(B, -1).cpu(), rec.ctx_slide) < tol, f"level {l}: slide context"
        if l < L - 1:
            # selection validity against the FREE-running oracle is only meaningful while the trajectories coincide;
            # check against the teacher-forced record (same inputs as ours at this level)
            for j in range(B):
                n = int(rec.num_ims[j])
                sc = rec.importance[j, :n]
                mine = forced[l][j]
                k = len(mine)
                if k < n:
                    kth = torch.sort(sc, descending=True).values[k - 1]
                    must = set(torch.nonzero(sc > kth + tol).flatten().tolist())
                    may = set(torch.nonzero(sc >= kth - tol).flatten().tolist())
                    assert must <= set(mine.tolist()) <= may, f"level {l} slide {j}: not a valid top-{k}"
                    s_m = sc[mine]
                    assert bool((s_m[:-1] >= s_m[1:] - tol).all()), "selection must be importance-descending"
                else:
                    assert sorted(mine.tolist()) == list(range(n))
    assert rel_err(hz.detach().cpu(), ohz) < tol
    assert abs(float(loss) - float(oloss)) < tol * max(1.0, abs(float(oloss)))
    # 3. gradients on every live parameter; None / zero pattern
    #   fp32   : every tensor within 1e-4 of its own scale (floored at 1e-3 of the largest gradient entry; tensors of a
    #            handful of elements -- the importance head's output bias -- are pure cancelling sums: floor 1e-2)
    #   bf16x3 : the whole flat gradient within 1e-4 (max-norm and L2) of the gradient's scale; single tensors, whose entries
    #            are strongly cancelling sums over patches, within 1e-2 of their own scale (split-bf16 products carry 2^-17)
    #   bf16   : ReLU / top-K masks may flip for near-zero pre-activations; relative L2 error of the flat gradient < 5e-2
    named = dict(model.named_parameters())
    gmax = max(float(g.abs().max()) for g in ograds.values() if g is not None)
    num = den = 0.0
    worst_abs = 0.0
    for k, og in ograds.items():
        mine = named[k].grad
        if og is None:
            assert mine is None or float(mine.abs().max()) == 0.0, f"{k}: reference leaves this gradient empty"
            continue
        assert mine is not None, k
        diff = (mine.detach().cpu().double() - og.double())
        num += float((diff ** 2).sum()); den += float((og.double() ** 2).sum())
        worst_abs = max(worst_abs, float(diff.abs().max()))
        if prec < 2:
            floor = (1e-3 if og.numel() > 8 else 1e-2) * gmax
            err = float(diff.abs().max()) / max(float(og.abs().max()), floor)
            assert err < (tol if prec == 0 else 1e-2), f"{k}: gradient error {err:.2e}"
    gl2 = (num / den) ** 0.5
    assert gl2 < (5e-2 if prec == 2 else 1e-4), f"flat gradient: relative L2 error {gl2:.2e}"
    if prec < 2:
        assert worst_abs / gmax < 1e-4, f"flat gradient: max-norm error {worst_abs / gmax:.2e}"
    # 4. the FREE-running oracle picks the same patches as we do, unless its own scores are tied within the tolerance
    #    at the selection boundary (torch.topk leaves tie order unspecified; we break ties towards the lowest row index)
    for j in range(B):
        for l in range(1, L):
            nf, nt = int(free[l].num_ims[j]), int(recs[l].num_ims[j])
            if nf == nt and torch.equal(free[l].locs[j, :nf], recs[l].locs[j, :nt]):
                continue
            n = int(free[l - 1].num_ims[j])
            sc = torch.sort(free[l - 1].importance[j, :n], descending=True).values
            k = ocfg.top_k_patches[l - 1]
            assert 0 <= k < n, f"slide {j} level {l}: selections differ although every row was kept"
            near_tie = bool(((sc[:-1] - sc[1:]).abs()[max(0, k - 1):k + 1] <= tol).any()) or bool((sc[:k][:-1] - sc[:k][1:] <= tol).any())
            assert near_tie, f"slide {j} level {l}: selected indices differ from the oracle without a near-tie"
            break


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
