"""Slide-level data parallelism on the CPU (gloo, world size 2): bucket layout of the flat gradient buffer, backward-order
bucketed all-reduce, and N-rank gradients == 1-rank gradients on the concatenated batch.  The
per-rank gradients come from the CPU oracle (the CUDA engine cannot run here); what is under test is the host-side
plumbing of paths_b200/parallel.py."""
import os
import socket

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import paths_oracle as po
from tests.common import load_fixture, fixture_inputs, product_config, rel_err


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def _layout(fx, sd):
    from paths_b200.engine import ParamLayout, make_dims
    cfg = product_config(fx)
    model = cfg.get_model()
    model.load_state_dict(sd, strict=True)
    return cfg, model, ParamLayout(model, make_dims(cfg.model_config, cfg))


def test_bucket_layout():
    from paths_b200.parallel import GradBuckets, shard_slides
    fx = load_fixture("small_full")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg, model, layout = _layout(fx, sd)
    bk = GradBuckets.from_layout(layout)
    assert bk.covers_exactly_once()
    names = [n for n, _, _ in bk.ranges]
    assert names == [f"procs.{l}" for l in reversed(range(cfg.num_levels))] + ["lstm"], names      # backward order
    # every live parameter sits inside the bucket of its own level
    for n, prm, off, k in layout.entries:
        owner = "lstm" if n.startswith("lstm.") else ".".join(n.split(".")[:2])
        a, b = next((a, b) for name, a, b in bk.ranges if name == owner)
        assert a <= off and off + k <= b, n
    assert shard_slides(5, 0, 2) == [0, 1, 2] and shard_slides(5, 1, 2) == [3, 4]
    big = shard_slides(4, 0, 2, weights=[20000, 50, 60, 70])
    assert big == [0] and shard_slides(4, 1, 2, weights=[20000, 50, 60, 70]) == [1, 2, 3]


def _worker(rank, world, port, out):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.set_num_threads(2)
    from paths_b200.parallel import GradBuckets, shard_slides
    fx = load_fixture("small_full")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg, model, layout = _layout(fx, sd)
    mine = shard_slides(len(slides), rank, world)
    _, loss, grads, _ = po.grads(sd, ocfg, [slides[i].grids for i in mine], ybin[mine], cens[mine])
    share = len(mine) / len(slides)                       # local mean -> share of the global mean
    layout.gflat.zero_()
    for n, prm, off, k in layout.entries:
        if grads[n] is not None:
            layout.gviews[n].copy_(grads[n] * share)
    bk = GradBuckets.from_layout(layout)
    bk.reduce_all(layout.gflat)
    gl = loss.clone() * share
    dist.all_reduce(gl)
    if rank == 0:
        torch.save({"gflat": layout.gflat.clone(), "loss": gl}, out)
    dist.barrier()
    dist.destroy_process_group()


@pytest.mark.timeout(300)
def test_two_rank_gradients_equal_single_rank(tmp_path):
    out = str(tmp_path / "dp.pt")
    mp.spawn(_worker, args=(2, _free_port(), out), nprocs=2, join=True)
    got = torch.load(out)
    fx = load_fixture("small_full")
    ocfg, sd, slides, ybin, cens = fixture_inputs(fx)
    cfg, model, layout = _layout(fx, sd)
    _, loss, grads, _ = po.grads(sd, ocfg, [s.grids for s in slides], ybin, cens)
    assert abs(float(got["loss"]) - float(loss)) < 1e-5 * max(1.0, abs(float(loss)))
    gmax = max(float(g.abs().max()) for g in grads.values() if g is not None)
    for n, prm, off, k in layout.entries:
        mine = got["gflat"][off:off + k].view(prm.shape)
        if grads[n] is None:
            assert float(mine.abs().max()) == 0.0, n
        else:
            assert float((mine - grads[n]).abs().max()) / max(float(grads[n].abs().max()), 1e-3 * gmax) < 1e-4, n
