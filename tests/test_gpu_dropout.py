"""Training-mode dropout of the aggregators (nn.Transformer(dropout=p), model/aggregator.py:25-33; live under model.train(),
train.py:131).  torch's generator cannot be reproduced, so the CUDA path hashes its keep/drop decisions (csrc/common.cuh);
the oracle rebuilds the same masks from (p, seed) and its mask placement is pinned against torch in
tests/test_oracle_golden.py::test_oracle_dropout_sites_match_torch.  With the masks shared, parity is exact again:
1e-4 in the fp32 and bf16x3 modes, the stated 2e-2 in the bf16 mode."""
import pytest
import torch

from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, product_config, product_model, product_batch, rel_err
from tests.parity_check import check_against_oracle

pytestmark = pytest.mark.gpu


def _step(name, prec, p, train=True, seed=11, slide_specs=None):
    from paths_b200.utils import inference_end2end
    fx = load_fixture(name)
    if slide_specs is not None:
        fx = dict(fx, slide_specs=slide_specs)
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg = product_config(fx, dropout=p)
    model = product_model(cfg, sd)
    model.train(train)
    rt = model.runtime()
    rt.set_precision(prec)
    batch, _ = product_batch(cfg, model, slides)
    torch.manual_seed(seed)
    hz, loss = inference_end2end(cfg.num_levels, cfg.top_k_patches, model, cfg.base_power, batch, cfg.task)
    loss.backward()
    torch.cuda.synchronize()
    return ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss


@pytest.mark.parametrize("prec", [0, 1, 2])
@pytest.mark.parametrize("name,p", [("small_full", 0.25), ("brca_b4", 0.05), ("small_rnn", 0.1)])
def test_dropout_step_matches_oracle_with_the_same_masks(name, p, prec):
    ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss = _step(name, prec, p)
    got_p, seed = rt.last_dropout
    assert got_p == pytest.approx(p) and seed != 0
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec, drop=po.HashDropout(p, seed))


@pytest.mark.parametrize("prec", [0, 1])
def test_dropout_long_bags(prec):
    """Bags beyond 128 tokens (level 0 of a 17 x 12 grid; level 2 keeps every row) take the tiled attention kernels, the short
    slide beside them the short-bag kernels of the same launch."""
    ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss = _step("small_full", prec, 0.2, slide_specs=[(0, 17, 12), (1, 4, 3), (2, 9, 9)])
    eng = rt.last_engine
    assert int(eng.buffer(0, "SEG")[1]) > 128
    check_against_oracle(ocfg, sd, slides, ybin, cens, cfg, model, rt, hz, loss, prec, drop=po.HashDropout(0.2, rt.last_dropout[1]))


def test_dropout_is_live_in_train_mode_only_and_seeded_by_torch():
    base = _step("brca_b4", 1, 0.0)
    ev = _step("brca_b4", 1, 0.2, train=False)
    assert ev[7].last_dropout == (0.0, 0)
    assert torch.equal(ev[8], base[8])                                   # model.eval(): dropout off, bit-identical hazards
    a = _step("brca_b4", 1, 0.2, seed=3)
    b = _step("brca_b4", 1, 0.2, seed=3)
    c = _step("brca_b4", 1, 0.2, seed=4)
    assert a[7].last_dropout == b[7].last_dropout != c[7].last_dropout   # seeds come from torch's CPU generator
    assert torch.equal(a[8], b[8]) and float(a[9]) == float(b[9])
    ga = {n: q.grad for n, q in a[6].named_parameters() if q.grad is not None}
    gb = {n: q.grad for n, q in b[6].named_parameters() if q.grad is not None}
    worst = max((rel_err(ga[n], gb[n]), n) for n in ga)                   # same masks: only the order of atomic bias sums differs
    assert worst[0] < 1e-5, worst
    assert rel_err(a[8], base[8]) > 1e-3 and rel_err(a[8], c[8]) > 1e-3  # dropout changes the step, and per seed


def test_dropout_keep_rate_on_the_device():
    """The FFN hidden activations of a level after the inner dropout: survivors are the positive entries that were kept, so their
    count falls to (1 - p) of the count without dropout (the upstream sites move individual entries across zero, not the rate),
    and every survivor carries the 1 / (1 - p) scale (the smallest positive value grows accordingly)."""
    p = 0.3
    base = _step("brca_b4", 0, 0.0)
    drop = _step("brca_b4", 0, p)
    eng0, eng1 = base[7].last_engine, drop[7].last_engine
    n = int(eng0.buffer(0, "NTOK").item()) * 512
    h0 = eng0.buffer(0, "HF0").float().flatten()[:n]
    h1 = eng1.buffer(0, "HF0").float().flatten()[:n]
    assert abs(float((h1 > 0).sum()) / float((h0 > 0).sum()) - (1 - p)) < 0.03
    assert abs(float(h1[h1 > 0].mean()) / float(h0[h0 > 0].mean()) - 1 / (1 - p)) < 0.1
