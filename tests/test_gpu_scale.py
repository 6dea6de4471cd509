"""Size-independent properties at a batch the oracle cannot reach (256 BRCA-shaped slides, ~69 k packed rows per step, 11-14 k per level: the regime
where the contractions run as cta_group::2 pairs with full stream-K plans and the level sets overlap): slides are independent
problems (`utils.py:228-279` loops over them), so
  * a slide's hazards do not depend on which other slides share its batch, and
  * the gradient of the batch equals the sum of the gradients of its two halves (same 1 / global-batch loss scale),
whatever tiles, pairs, row groups and stream-K segments the rows fall into in the three runs."""
import copy

import pytest
import torch

from paths_b200 import synth
from tests.common import BASE_CONFIG

pytestmark = pytest.mark.gpu
B = 256


def _setup(prec):
    from paths_b200.config import Config
    from paths_b200.data_utils.slide import load_patch_preprocessed_slide
    from paths_b200.engine import PyramidStore
    data = copy.deepcopy(BASE_CONFIG)
    data["model_config"]["dropout"] = 0.0
    cfg = Config.from_dict(data, test_mode=True)
    dev = torch.device("cuda")
    torch.manual_seed(7)
    model = cfg.get_model().to(dev)
    model.train()
    rt = model.runtime()
    rt.set_precision(prec)
    Lv, D = cfg.num_levels, cfg.model_config.patch_embed_dim
    store = PyramidStore(Lv, D, dev)
    ids, ybin, cens = [], [], []
    for sid in range(B):
        s = synth.make_device_slide_fast(sid, *synth.brca_shape(sid), Lv, D, 0.3, dev)
        sl = load_patch_preprocessed_slide(f"/x/s{sid}.svs", cfg.base_power, cfg.model_config.patch_size, model.procs[0].ctx_dim(), Lv,
                                           grids=s.grids, store=store)
        ids.append(sl.store_index); ybin.append(s.survival_bin); cens.append(s.censored)
    return cfg, rt, store, ids, ybin, cens, dev


def _step(cfg, rt, store, ids, ybin, cens, dev, scale):
    caps, maxseg = store.capacities(ids, cfg.top_k_patches)
    eng = rt.engine_for(len(ids), caps, maxseg, cfg.top_k_patches)
    eng.bind_store(store)
    eng.forward(torch.tensor(ids, dtype=torch.int32, device=dev), torch.tensor(ybin, dtype=torch.int32, device=dev),
                torch.tensor(cens, dtype=torch.int32, device=dev), scale, True)
    over, need = eng.overflow_poll()
    assert not over, need
    eng.backward(zero_grads=True)
    torch.cuda.synchronize()
    n = len(ids)
    return eng.buffer(0, "HAZARDS").view(n, -1).clone(), rt.layout.gflat.clone(), eng.counts()


@pytest.mark.parametrize("prec", [1, 2])
def test_batch_gradient_is_the_sum_over_its_halves_at_256_slides(prec):
    cfg, rt, store, ids, ybin, cens, dev = _setup(prec)
    hz, g, counts = _step(cfg, rt, store, ids, ybin, cens, dev, 1.0 / B)
    assert min(counts) > 10000                                            # the pair / full-wave regime
    h = B // 2
    hz_a, g_a, _ = _step(cfg, rt, store, ids[:h], ybin[:h], cens[:h], dev, 1.0 / B)
    hz_b, g_b, _ = _step(cfg, rt, store, ids[h:], ybin[h:], cens[h:], dev, 1.0 / B)
    assert torch.isfinite(g).all() and float(g.abs().max()) > 0
    # a row's products do not depend on the tile it sits in; only reductions over rows (weight gradients) change their order
    assert float((torch.cat([hz_a, hz_b]) - hz).abs().max()) <= 1e-6 * float(hz.abs().max())
    gmax = float(g.abs().max())
    assert float((g_a + g_b - g).abs().max()) <= 2e-5 * gmax
    assert float((g_a + g_b - g).norm() / g.norm()) <= 1e-5
