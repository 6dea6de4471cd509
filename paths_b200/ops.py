"""torch.library custom ops over the C ABI (include/paths_b200.h): the "thin torch custom-op layer" between the Python mirror
of the reference's API and libpaths_b200.so.

    torch.ops.paths_b200.step(anchor, engine, slide_ids, fts_padded, fts_packed, ybin, censored, scale, graph) -> (hazards, loss)

is one whole hierarchical step -- utils.inference_end2end (utils.py:228-279): paths_engine_forward[_padded|_packed] -- with a
registered autograd formula whose backward is loss.backward() of train.py:65 (paths_engine_backward, gradients accumulated
straight into the flat buffer the parameters' .grad are views of; with an attached DataParallelStep also the bucketed
all-reduce).  `engine` is an integer handle into a registry of live (runtime, engine, graph-step) triples: custom ops carry
tensors and scalars only.  The op mutates engine-owned workspace that is not among its arguments, so it is registered without
functionalisation metadata and is meant for eager use (the engine sequences its own streams / CUDA graphs; there is no
torch.compile path, by design).
"""
from __future__ import annotations

import weakref
from typing import Dict, Optional, Tuple

import torch
from torch import Tensor

_ENGINES: Dict[int, tuple] = {}


def register_engine(rt, eng, gs=None) -> int:
    """Handle of a (runtime, engine, graph step) triple for the custom ops; entries die with their engine."""
    key = id(eng)
    ent = _ENGINES.get(key)
    if ent is None or ent[1]() is not eng:
        _ENGINES[key] = (weakref.ref(rt), weakref.ref(eng, lambda _r, k=key: _ENGINES.pop(k, None)), [None])
    _ENGINES[key][2][0] = gs
    return key


def _lookup(key: int):
    rt_ref, eng_ref, gs = _ENGINES[key]
    rt, eng = rt_ref(), eng_ref()
    if rt is None or eng is None:
        raise RuntimeError("paths_b200::step: the engine behind this handle no longer exists")
    return rt, eng, gs[0]


@torch.library.custom_op("paths_b200::step", mutates_args=())
def step(anchor: Tensor, engine: int, slide_ids: Tensor, fts_padded: Optional[Tensor], fts_packed: Optional[Tensor], ybin: Tensor,
         censored: Optional[Tensor], scale: float, graph: bool) -> Tuple[Tensor, Tensor]:
    rt, eng, gs = _lookup(engine)
    if graph and gs is not None:                            # CUDA-graph replay of the same enqueue (runtime.GraphStep)
        gs.load(slide_ids, ybin, censored, fts_packed)
        gs.g_fwd.replay()
    else:
        eng.forward(slide_ids, ybin, censored, scale, True, fts_padded, fts_packed)
    hz = eng.buffer(0, "HAZARDS").view(eng.B, eng.dims.nl).clone()
    loss = eng.buffer(0, "LOSS").view(()).clone()
    return hz, loss


@step.register_fake
def _(anchor, engine, slide_ids, fts_padded, fts_packed, ybin, censored, scale, graph):
    rt, eng, _gs = _lookup(engine)
    return anchor.new_empty((eng.B, eng.dims.nl)), anchor.new_empty(())


def _setup(ctx, inputs, output):
    ctx.engine, ctx.graph = inputs[1], inputs[8]
    ctx.save_for_backward(output[0])


def _backward(ctx, g_hz, g_loss):
    rt, eng, gs = _lookup(ctx.engine)
    (hz,) = ctx.saved_tensors
    dlog = eng.buffer(eng.dims.L - 1, "DLOGITS").view(eng.B, eng.dims.nl)
    if g_loss is not None:
        dlog.mul_(g_loss)
    else:
        dlog.zero_()
    if g_hz is not None and rt.track_hazard_grad:
        dlog.add_(g_hz * hz * (1 - hz) if eng.dims.task == 0 else g_hz)
    rt.layout.prepare_grads()
    if rt.grad_sync is not None and rt.grad_sync.world > 1:  # data parallelism: bucketed all-reduce behind the backward pass
        rt.grad_sync.backward(eng, zero=False)
    elif ctx.graph and gs is not None and gs.g_bwd is not None:
        gs.g_bwd.replay()
        rt.layout.attach_grads()
    else:
        eng.backward(False)
        rt.layout.attach_grads()
    return (None,) * 9


step.register_autograd(_backward, setup_context=_setup)
