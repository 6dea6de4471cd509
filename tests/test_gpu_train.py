"""GPU tests of the pieces around the step (e): fused flat AdamW against torch.optim.AdamW, checkpoint
round-trip with the reference's key names, and the level-by-level data-parallel backward against the monolithic one."""
import copy

import pytest
import torch

from tests.common import load_fixture, fixture_inputs, product_config, product_model, product_batch, rel_err

pytestmark = pytest.mark.gpu


def _setup(name="small_full"):
    fx = load_fixture(name)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    batch, store = product_batch(cfg, model, slides)
    return cfg, model, batch


def test_flat_adamw_matches_torch_adamw():
    from paths_b200.optim import FlatAdamW
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup()
    opt = FlatAdamW(model, lr=1e-3, weight_decay=1e-2)
    sched = cfg.get_lr_scheduler(opt)                                     # ExponentialLR must accept it (config.py:130)
    ref = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters()}
    ropt = torch.optim.AdamW(list(ref.values()), lr=1e-3, weight_decay=1e-2)
    rsched = torch.optim.lr_scheduler.ExponentialLR(ropt, cfg.lr_decay_per_epoch)
    for step in range(4):
        opt.zero_grad()
        ropt.zero_grad()
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
        for n, p in model.named_parameters():
            ref[n].grad = None if p.grad is None else p.grad.detach().clone()
        opt.step()
        ropt.step()
        if step == 1:
            sched.step(); rsched.step()
        for n, p in model.named_parameters():
            assert rel_err(p.detach(), ref[n].detach()) < 2e-6, (step, n)
    none_grad = [n for n, p in model.named_parameters() if p.grad is None]
    assert any("classification_layer" in n for n in none_grad)            # untouched, exactly as torch leaves them
    assert abs(opt.param_groups[0]["lr"] - ropt.param_groups[0]["lr"]) < 1e-12


def test_flat_adamw_leaves_never_evaluated_modules_alone():
    """RNN variant: procs.0.hctx_mlp is never evaluated (paths.py:102), its .grad stays None in the reference and
    torch.optim.AdamW skips it -- no weight decay either; the dead encoder weights (exact-zero gradients) are decayed."""
    from paths_b200.optim import FlatAdamW
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("small_rnn")
    before = {n: p.detach().clone() for n, p in model.named_parameters()}
    opt = FlatAdamW(model, lr=1e-2, weight_decay=0.5)
    ref = {n: p.detach().clone().requires_grad_(True) for n, p in model.named_parameters()}
    ropt = torch.optim.AdamW(list(ref.values()), lr=1e-2, weight_decay=0.5)
    for _ in range(2):
        opt.zero_grad(); ropt.zero_grad()
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
        for n, p in model.named_parameters():
            ref[n].grad = None if p.grad is None else p.grad.detach().clone()
        opt.step(); ropt.step()
    named = dict(model.named_parameters())
    skipped = [n for n, p in named.items() if p.grad is None]
    assert any(n.startswith("procs.0.hctx_mlp") for n in skipped)
    for n in skipped:
        assert torch.equal(named[n].detach(), before[n]), n
    enc = "procs.1.global_agg.transformer.encoder.layers.0.linear1.weight"
    assert not torch.equal(named[enc].detach(), before[enc])
    for n, p in named.items():
        assert rel_err(p.detach(), ref[n].detach()) < 2e-6, n


def test_checkpoint_round_trip(tmp_path):
    from paths_b200.utils import save_state, load_state, inference_end2end
    cfg, model, batch = _setup()
    with torch.no_grad():
        hz0, loss0 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    save_state(str(tmp_path), model, {"epoch": 7, "train_loss": {1: 0.5}})
    sd = torch.load(str(tmp_path / "model.pt"))
    assert set(sd.keys()) == set(model.state_dict().keys())
    with torch.no_grad():
        for p in model.parameters():
            p.mul_(0.5)
        hz1, _ = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    assert rel_err(hz1, hz0) > 1e-3
    stats = load_state(str(tmp_path), model)
    assert stats["epoch"] == 7
    with torch.no_grad():
        hz2, loss2 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    assert torch.equal(hz2, hz0) and torch.equal(loss2, loss0)


def test_dp_step_equals_autograd_step():
    """DataParallelStep (level-by-level backward, buckets) on one rank == loss.backward() through the autograd node."""
    from paths_b200.parallel import DataParallelStep
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    want = {n: (None if p.grad is None else p.grad.detach().clone()) for n, p in model.named_parameters()}
    for p in model.parameters():
        p.grad = None
    dp = DataParallelStep(model)
    hz2, loss2 = dp.step(copy.copy(batch), cfg.top_k_patches)
    assert torch.equal(hz2, hz.detach()) and torch.equal(loss2, loss.detach())
    for n, p in model.named_parameters():
        if want[n] is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
        else:
            # weight gradients accumulate with fp32 atomics: order may differ between runs
            assert rel_err(p.grad, want[n]) < 1e-5, n


def test_eval_batch_after_a_dp_step_is_a_mean_over_its_own_slides():
    """A data-parallel training step scales the loss by 1 / global batch; that scale must not leak into the evaluation
    batches that follow it (train.py:78-83: validation through inference_end2end after the epoch's steps)."""
    from paths_b200.parallel import DataParallelStep
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    with torch.no_grad():
        _, want = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, copy.copy(batch), cfg.task)
    dp = DataParallelStep(model)
    _, part = dp.step(copy.copy(batch), cfg.top_k_patches, global_batch=64)           # this rank's share of a 64-slide batch
    assert abs(float(part) * 64 / len(batch["slide"]) - float(want)) < 1e-6 * abs(float(want)) + 1e-7
    with torch.no_grad():
        _, again = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, copy.copy(batch), cfg.task)
    assert torch.equal(again, want)


def test_packed_level0_rows_equal_the_padded_batch():
    """collate_fn(packed=True): the step fed with the unpadded level-0 rows computes bit-identical results to the padded batch,
    directly and through the prefetcher (which then leaves the padded copy on the host)."""
    from paths_b200.data_utils.dataset import collate_fn
    from paths_b200.data_utils.prefetch import DevicePrefetcher
    from paths_b200.data_utils.slide import load_patch_preprocessed_slide
    from paths_b200.utils import inference_end2end
    fx = load_fixture("brca_b4")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    batch, store = product_batch(cfg, model, slides)
    items = [sl.todict() | {"survival_bin": s.survival_bin, "survival": s.survival, "censored": s.censored, "slide": sl}
             for s, sl in zip(slides, batch["slide"])]
    pbatch = collate_fn(items, packed=True)
    assert pbatch["fts_packed"].shape[0] == int(pbatch["num_ims"].sum()) and torch.equal(pbatch["fts"], batch["fts"])
    with torch.no_grad():
        hz0, loss0 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        hz1, loss1 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, pbatch, cfg.task)
        assert torch.equal(hz0, hz1) and torch.equal(loss0, loss1)
        for b in DevicePrefetcher([pbatch], "cuda"):
            assert b["fts_packed"].is_cuda and not b["fts"].is_cuda
            hz2, loss2 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
            assert torch.equal(hz0, hz2) and torch.equal(loss0, loss2)


def test_attached_dp_step_runs_inside_loss_backward():
    """DataParallelStep.attach(): the reference's loop body (inference_end2end + loss.backward()) gives the gradients of
    dp.step(); detach() restores the plain path."""
    from paths_b200.parallel import DataParallelStep
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    dp = DataParallelStep(model)
    hz, loss = dp.step(copy.copy(batch), cfg.top_k_patches, global_batch=8)
    want = {n: (None if p.grad is None else p.grad.detach().clone()) for n, p in model.named_parameters()}
    for p in model.parameters():
        p.grad = None
    dp.attach(global_batch=8)
    hz2, loss2 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, copy.copy(batch), cfg.task)
    loss2.backward()
    dp.detach()
    assert model.runtime().grad_sync is None
    assert torch.equal(hz2.detach(), hz)
    for n, p in model.named_parameters():
        if want[n] is None:
            assert p.grad is None or float(p.grad.abs().max()) == 0.0, n
        else:
            assert rel_err(p.grad, want[n]) < 1e-5, n


def test_bfloat16_pyramid_store():
    """PyramidStore(dtype=bfloat16): half the footprint; the step equals the oracle fed with the features rounded once."""
    from oracle import paths_oracle as po
    from paths_b200.engine import PyramidStore
    from paths_b200.utils import inference_end2end
    from tests.parity_check import check_against_oracle
    fx = load_fixture("brca_b4")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)
    model = product_model(cfg, sd)
    rt = model.runtime(); rt.set_precision(0)
    store = PyramidStore(cfg.num_levels, cfg.model_config.patch_embed_dim, "cuda", dtype=torch.bfloat16)
    batch, _ = product_batch(cfg, model, slides, store=store)
    assert store.grids[0][2].dtype == torch.bfloat16
    for use_fts in (True, False):                                   # level-0 rows from the batch / gathered from the store
        b = dict(batch)
        if not use_fts:
            b.pop("fts")
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        for p in model.parameters():
            p.grad = None
        loss.backward()
        torch.cuda.synchronize()
        for s in slides:
            s.grids = [g.to(torch.bfloat16).float() for g in s.grids]
        check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, 0)


def test_custom_op_layer_equals_the_autograd_function():
    """torch.ops.paths_b200.step (paths_b200/ops.py) is the default route of inference_end2end; the plain autograd.Function
    route (PATHS_B200_NO_CUSTOM_OP=1) must give the same hazards, loss and gradients."""
    from paths_b200.utils import inference_end2end
    res = []
    for custom in (True, False):
        cfg, model, batch = _setup("brca_b4")
        rt = model.runtime(); rt.use_custom_op = custom
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        assert loss.requires_grad
        (2.0 * loss).backward()
        res.append((hz.detach().clone(), loss.detach().clone(), rt.layout.gflat.clone()))
    assert torch.ops.paths_b200.step is not None
    assert torch.equal(res[0][0], res[1][0]) and torch.equal(res[0][1], res[1][1])
    assert rel_err(res[0][2], res[1][2]) < 1e-5


def test_graph_replay_equals_eager_steps():
    """rt.use_graphs: the step replayed as CUDA graphs (forward graph + backward graph, inputs in static buffers) gives the
    eager step's hazards, loss and gradients -- on the batch it was captured with and on a different batch of the same size;
    gradients accumulate across backward calls as autograd's do."""
    from paths_b200.data_utils.dataset import collate_fn
    from paths_b200.utils import inference_end2end
    fx = load_fixture("brca_b4")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx)

    def batches(model):
        from tests.common import product_batch as pb
        b1, store = pb(cfg, model, slides)
        order = [2, 0, 3, 1]
        b2, _ = pb(cfg, model, [slides[i] for i in order], store=store)
        out = []
        for b in (b1, b2):
            items = [sl.todict() | {"survival_bin": int(b["survival_bin"][i]), "survival": float(b["survival"][i]),
                                    "censored": int(b["censored"][i]), "slide": sl} for i, sl in enumerate(b["slide"])]
            out.append(collate_fn(items, packed=True))
        return out

    res = {}
    for mode in ("eager", "graph"):
        model = product_model(cfg, sd)
        rt = model.runtime(); rt.set_precision(1); rt.use_graphs = mode == "graph"
        got = []
        for b in batches(model) * 2:
            for p in model.parameters():
                p.grad = None
            hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
            loss.backward()
            got.append((hz.detach().clone(), loss.detach().clone(), rt.layout.gflat.clone()))
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        loss.backward()                                                   # no zero_grad in between: accumulates
        got.append((hz.detach().clone(), loss.detach().clone(), rt.layout.gflat.clone()))
        if mode == "graph":
            assert len(rt._graphs) == 1
        res[mode] = got
    for (h0, l0, g0), (h1, l1, g1) in zip(res["eager"], res["graph"]):
        assert torch.equal(h0, h1) and torch.equal(l0, l1)
        assert rel_err(g1, g0) < 1e-5
    assert rel_err(res["graph"][4][2], 2 * res["graph"][3][2]) < 1e-5


def test_device_prefetcher_same_results_as_host_batches():
    """Batches staged one step ahead by DevicePrefetcher give bit-identical hazards/loss to host batches handed over directly."""
    from paths_b200.data_utils.prefetch import DevicePrefetcher
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    with torch.no_grad():
        hz0, loss0 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    n = 0
    for b in DevicePrefetcher((batch for _ in range(3)), "cuda"):
        assert b["fts"].is_cuda and not batch["fts"].is_cuda            # the host batch itself is left untouched
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, b, cfg.task)
        loss.backward()
        assert torch.equal(hz.detach(), hz0) and torch.equal(loss.detach(), loss0)
        n += 1
    assert n == 3


def test_engines_sharing_parameters_never_use_stale_packed_weights():
    """The step engine and a level-API engine are bound to the same parameters and share the packed-weight registry.  After
    the level engine took the registry over and the weights changed, the step engine must re-pack and contract against its own
    copy: its result equals that of a fresh model built from the updated weights."""
    from paths_b200.data_utils.patch_batch import from_batch
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    keys = ("fts", "locs", "parent_inds", "ctx_patch", "ctx_slide", "num_ims")
    with torch.no_grad():
        inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)          # step engine packs
        model(0, from_batch({k: batch[k].clone() for k in keys}, "cuda"))                                       # level engine registers
        for p in model.parameters():
            p.mul_(0.9)
        hz1, loss1 = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        out_level = model(0, from_batch({k: batch[k].clone() for k in keys}, "cuda"))
        sd = {k: v.detach().cpu().clone() for k, v in model.state_dict().items()}
    fx = load_fixture("brca_b4")
    fresh = product_model(product_config(fx), sd)
    with torch.no_grad():
        batch2, _ = product_batch(cfg, fresh, fixture_inputs(fx)[2])
        hz2, loss2 = inference_end2end(cfg.num_levels, cfg.top_k_patches, fresh, cfg.base_power, batch2, cfg.task)
        out_level2 = fresh(0, from_batch({k: batch2[k].clone() for k in keys}, "cuda"))
    assert torch.equal(hz1, hz2) and torch.equal(loss1, loss2)
    assert torch.equal(out_level["importance"], out_level2["importance"])


def test_step_is_reproducible():
    """Level sets run concurrently on engine-owned streams: repeated steps on the same batch must give the same loss bit for
    bit (forward has no atomics) and gradients equal up to the order of the fp32 atomics of the LayerNorm / bias sums."""
    from paths_b200.utils import inference_end2end
    cfg, model, batch = _setup("brca_b4")
    ref_loss, ref_g = None, None
    for it in range(8):
        for p in model.parameters():
            p.grad = None
        hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
        loss.backward()
        torch.cuda.synchronize()
        g = model.runtime().layout.gflat.clone()
        if ref_loss is None:
            ref_loss, ref_g = loss.detach().clone(), g
        else:
            assert torch.equal(loss.detach(), ref_loss), it
            assert rel_err(g, ref_g) < 1e-5, it


def test_device_prefetcher_early_exit_and_loader_errors():
    """The helper thread of DevicePrefetcher: leaving the loop early shuts it down, an exception of the wrapped loader
    surfaces in the consuming loop, device placement and values are those of the host batches."""
    from paths_b200.data_utils.prefetch import DevicePrefetcher

    def mk(i):
        return {"fts": torch.full((2, 3, 4), float(i)), "survival_bin": torch.tensor([i, i + 1]), "censored": torch.tensor([0, 1]),
                "slide": ["a", "b"]}

    it = iter(DevicePrefetcher((mk(i) for i in range(5)), "cuda"))
    b0 = next(it)
    assert b0["fts"].is_cuda and b0["survival_bin"].is_cuda and b0["slide"] == ["a", "b"]
    assert float(b0["fts"].sum()) == 0.0
    b1 = next(it)
    assert float(b1["fts"].mean()) == 1.0
    it.close()                                                  # early exit: the worker is joined, nothing hangs

    def bad():
        yield mk(0)
        raise ValueError("loader failed")

    with pytest.raises(ValueError, match="loader failed"):
        for _ in DevicePrefetcher(bad(), torch.device("cuda", 0)):
            pass
    assert [float(b["fts"].mean()) for b in DevicePrefetcher((mk(i) for i in range(4)), "cuda")] == [0.0, 1.0, 2.0, 3.0]
