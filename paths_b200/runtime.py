"""Binds a RecursiveModel's parameters to level engines and exposes the two reference-facing operations:
the whole hierarchical step (utils.inference_end2end) with autograd, and one level on a padded PatchBatch
(RecursiveModel.forward)."""
from __future__ import annotations

import os
import warnings
from typing import Dict, List, Optional, Sequence

import torch
import torch.distributed

from . import _lib as L
from .engine import Engine, ParamLayout, PyramidStore, make_dims, pe_div_table

PRECISIONS = {"fp32": 0, "bf16x3": 1, "bf16": 2}


def default_precision() -> int:
    return PRECISIONS[os.environ.get("PATHS_B200_PRECISION", "bf16x3")]


class _StepFn(torch.autograd.Function):
    """Autograd node of one whole hierarchical step.  Gradients of the parameters are accumulated by the engine
    directly into the flat gradient buffer the parameters' .grad are views of (no per-parameter copies)."""

    @staticmethod
    def forward(ctx, anchor, rt, eng, slide_ids, fts_padded, ybin, cens, scale, fts_packed=None, gs=None):
        if gs is not None:                                  # CUDA-graph replay of the same enqueue (GraphStep)
            gs.load(slide_ids, ybin, cens, fts_packed)
            gs.g_fwd.replay()
        else:
            eng.forward(slide_ids, ybin, cens, scale, True, fts_padded, fts_packed)   # whole step: chain + grouped aggregators on the engine's streams
        ctx.rt, ctx.eng, ctx.gs = rt, eng, gs
        hz = eng.buffer(0, "HAZARDS").view(eng.B, eng.dims.nl).clone()
        loss = eng.buffer(0, "LOSS").view(()).clone()
        ctx.save_for_backward(hz)
        return hz, loss

    @staticmethod
    def backward(ctx, g_hz, g_loss):
        rt, eng = ctx.rt, ctx.eng
        (hz,) = ctx.saved_tensors
        dlog = eng.buffer(eng.dims.L - 1, "DLOGITS").view(eng.B, eng.dims.nl)
        if g_loss is not None:
            dlog.mul_(g_loss)
        else:
            dlog.zero_()
        if g_hz is not None and rt.track_hazard_grad:
            dlog.add_(g_hz * hz * (1 - hz) if eng.dims.task == 0 else g_hz)
        rt.layout.prepare_grads()
        if rt.grad_sync is not None and rt.grad_sync.world > 1:   # data parallelism: bucketed all-reduce behind the backward pass
            rt.grad_sync.backward(eng, zero=False)
        elif ctx.gs is not None and ctx.gs.g_bwd is not None:
            ctx.gs.g_bwd.replay()
            rt.layout.attach_grads()
        else:
            eng.backward(False)
            rt.layout.attach_grads()
        return (None,) * 10


class ModelRuntime:
    def __init__(self, model, precision: Optional[int] = None):
        self.model = model
        self.precision = default_precision() if precision is None else precision
        self.dims = make_dims(model.model_config, model.train_config, self.precision)
        first = next(model.parameters())
        if not first.is_cuda:
            raise RuntimeError("paths_b200: move the model to a CUDA device first (model.to('cuda')); there is no CPU path")
        self.device = first.device
        self.layout = ParamLayout(model, self.dims)
        self.div = pe_div_table(self.dims, self.device)
        self.anchor = torch.zeros((), device=self.device, requires_grad=True)
        self.engines: Dict[tuple, Engine] = {}
        self.track_hazard_grad = False
        self.overflow_reruns = 0                        # steps re-run with worst-case buffers (no-child fallback, slide.py:336-352)
        self.grad_sync = None                           # parallel.DataParallelStep.attach(): loss.backward() also all-reduces
        # the step as a torch.library custom op (paths_b200::step) or, with PATHS_B200_NO_CUSTOM_OP=1, as a plain autograd.Function
        self.use_custom_op = os.environ.get("PATHS_B200_NO_CUSTOM_OP", "0") != "1"
        # CUDA-graph replay of the step (GraphStep) for host batches that carry packed level-0 rows or read the resident store:
        # the small-batch regime, where ~270 launches per step outrun the host.  Off by default (a graph pins its engine's buffers).
        self.use_graphs = os.environ.get("PATHS_B200_GRAPHS", "0") == "1"
        self._graphs: Dict[tuple, "GraphStep"] = {}
        self.last_dropout = (0.0, 0)                    # (p, seed) of the latest forward pass (tests rebuild the masks from it)
        self._ids_cache: Dict[tuple, torch.Tensor] = {}

    # ------------------------------------------------------------------------------------------------------------
    def set_precision(self, precision) -> None:
        self.precision = PRECISIONS[precision] if isinstance(precision, str) else int(precision)
        self.dims.precision = self.precision
        for e in self.engines.values():
            e.set_precision(self.precision)
        self._graphs.clear()

    def _dims_for(self, top_k: Sequence[int]) -> L.Dims:
        d = L.Dims.from_buffer_copy(self.dims)
        for i in range(L.MAX_LEVELS):
            d.K[i] = int(top_k[i]) if i < len(top_k) else -1
        d.precision = self.precision
        return d

    def engine_for(self, B: int, caps: Sequence[int], maxseg: Sequence[int], top_k: Sequence[int], tag: str = "step") -> Engine:
        key = (tag, B, tuple(int(k) for k in top_k))
        eng = self.engines.get(key)
        if eng is None or not eng.fits(B, caps, maxseg):
            if eng is not None:
                caps = [max(a, b) for a, b in zip(caps, eng.caps)]
                maxseg = [max(a, b) for a, b in zip(maxseg, eng.maxseg)]
                del self.engines[key]
                self._graphs = {k: g for k, g in self._graphs.items() if g.eng is not eng}     # captured graphs pin its buffers
                del eng
                torch.cuda.empty_cache()
            eng = Engine(self._dims_for(top_k), B, caps, maxseg, self.device)
            eng.bind_params(self.layout, self.div)
            self.engines[key] = eng
        return eng

    def _arm_dropout(self, eng: Engine) -> None:
        """nn.Transformer's dropout is live in model.train() (train.py:131) and off in model.eval().  The keep/drop decisions
        are hashed from a per-step seed drawn from torch's CPU generator, so torch.manual_seed makes runs repeatable."""
        p = float(self.model.model_config.dropout) if self.model.training else 0.0
        seed = int(torch.randint(0, 2 ** 62, (1,), dtype=torch.int64).item()) if p > 0 else 0
        if p > 0 and torch.distributed.is_available() and torch.distributed.is_initialized():
            # data-parallel ranks seeded alike must not drop the same (row, column) positions of their shards
            seed = (seed + torch.distributed.get_rank() * 0x9E3779B97F4A7C15) & 0x3FFFFFFFFFFFFFFF
        eng.set_dropout(p, seed)
        self.last_dropout = (p, seed)

    # ------------------------------------------------------------------------------------------------------------
    def end2end(self, batch: Dict, keep_patches: Sequence[int], use_batch_fts: bool = True, global_batch: Optional[int] = None):
        """utils.inference_end2end (utils.py:228-279) on the device.  `batch` is the collated dict of
        data_utils/dataset.py:229-243 whose "slide" entries are paths_b200 PreprocessedSlide objects.  `global_batch` (data
        parallelism): the loss is this rank's share of the mean over that many slides instead of the mean over the batch.

        Level buffers are sized for the normal case (at most 4 children per kept row).  A slide that takes the reference's
        no-child fallback (slide.py:336-352: every non-background cell of the next grid) can exceed them: the kernels then raise
        a device flag, which is read here once the last select of the pass has run (the rest of the pass is still queued, so
        the wait costs no GPU idle time), and the step is re-run on an engine sized for the worst case BEFORE any output or
        autograd node of the truncated pass is handed out."""
        slides = batch["slide"]
        B = len(slides)
        store: PyramidStore = slides[0].store
        ids = [s.store_index for s in slides]
        top_k = list(keep_patches)
        worst = False
        while True:
            caps, maxseg = store.capacities(ids, top_k, worst_case=worst)
            eng = self.engine_for(B, caps, maxseg, top_k)
            if getattr(eng, "_store", (None,))[0] is not store or store.tables()[1] is not eng._store[2]:
                eng.bind_store(store)
            dev = self.device
            key = tuple(ids)
            slide_ids = self._ids_cache.get(key)               # the same batches recur every epoch: no per-step host -> device copy
            if slide_ids is None:
                if len(self._ids_cache) > 4096:
                    self._ids_cache.clear()
                slide_ids = self._ids_cache[key] = torch.tensor(ids, dtype=torch.int32).to(dev)
            # level-0 features: the batch's own rows -- packed ("fts_packed", our collate_fn's extra key: the same rows without
            # dataset.py:217's zero padding, 43 % fewer bytes across PCIe on BRCA-shaped bags) or the reference's padded "fts"
            fts = pk = None
            if use_batch_fts and batch.get("fts_packed") is not None:
                pk = batch["fts_packed"].to(dev, dtype=torch.float32, non_blocking=True).contiguous()
            elif use_batch_fts and batch.get("fts") is not None:
                fts = batch["fts"].to(dev, dtype=torch.float32, non_blocking=True).contiguous()
            task = self.dims.task
            if task == 0:
                ybin = torch.as_tensor(batch["survival_bin"]).to(dev, dtype=torch.int32, non_blocking=True)
                cens = torch.as_tensor(batch["censored"]).to(dev, dtype=torch.int32, non_blocking=True)
            else:
                ybin = torch.as_tensor(batch["subtype"]).to(dev, dtype=torch.int32, non_blocking=True)
                cens = None
            gb = global_batch
            if gb is None and self.grad_sync is not None and torch.is_grad_enabled():     # attached data-parallel training step
                gb = self.grad_sync.global_batch or B * self.grad_sync.world
            scale = 1.0 / float(gb if gb else B)
            self._arm_dropout(eng)
            gs = None
            if self.use_graphs and self.last_dropout[0] == 0.0 and fts is None and (self.grad_sync is None or self.grad_sync.world == 1):
                gs = self._graph_for(eng, slide_ids, ybin, cens, scale, pk, torch.is_grad_enabled())
            if self.use_custom_op:                      # torch.ops.paths_b200.step (ops.py): the custom-op layer over the C ABI
                from . import ops
                hz, loss = ops.step(self.anchor, ops.register_engine(self, eng, gs), slide_ids, fts, pk, ybin, cens, scale, gs is not None)
            elif torch.is_grad_enabled():
                hz, loss = _StepFn.apply(self.anchor, self, eng, slide_ids, fts, ybin, cens, scale, pk, gs)
            else:
                hz, loss = _StepFn.forward(_NoCtx(), self.anchor, self, eng, slide_ids, fts, ybin, cens, scale, pk, gs)
            over, need = (False, 0) if os.environ.get("PATHS_B200_SKIP_POLL") == "1" else eng.overflow_poll()   # (diagnosis only)
            if over:
                if worst:
                    raise RuntimeError(f"paths_b200: level buffers sized for the worst case still overflowed ({need} rows needed)")
                worst = True
                self.overflow_reruns += 1
                del hz, loss
                continue
            self.last_engine = eng
            return hz, loss

    def _graph_for(self, eng, slide_ids, ybin, cens, scale, pk, backward) -> "GraphStep":
        key = (id(eng), self.precision, float(scale), pk is not None, bool(backward), cens is not None)
        gs = self._graphs.get(key)
        if gs is None or gs.eng is not eng:
            fts = None
            if pk is not None:                          # static input buffer for the level-0 rows, sized for the engine's capacity
                fts = torch.zeros(eng.caps[0], pk.shape[1], device=self.device)
                fts[:pk.shape[0]].copy_(pk)
            gs = self._graphs[key] = GraphStep(self, eng, slide_ids, ybin, cens, scale, backward=False, fts_packed=fts)
            if backward:                                # the autograd node replays the backward pass without clearing the gradients
                self.layout.prepare_grads()
                st = torch.cuda.Stream(self.device)
                st.wait_stream(torch.cuda.current_stream(self.device))
                gs.g_bwd = torch.cuda.CUDAGraph()
                with torch.cuda.graph(gs.g_bwd, stream=st, capture_error_mode="thread_local"):
                    eng.backward(False)
                gs.has_backward = True
                self.layout.gflat.zero_()
        return gs

    @torch.no_grad()
    def aggregate(self, depth: int, tokens: torch.Tensor, lengths: Optional[torch.Tensor]) -> torch.Tensor:
        return _aggregate(self, depth, tokens, lengths)

    def graph_step(self, eng, slide_ids, ybin, cens, scale, backward=True, fts_packed=None) -> "GraphStep":
        """Capture one step of `eng` (forward, optionally backward) as CUDA graphs; see GraphStep."""
        return GraphStep(self, eng, slide_ids, ybin, cens, scale, backward, fts_packed)

    def risk(self) -> torch.Tensor:
        """Survival risk scores of the latest step, -sum(cumprod(1 - hazards)) per slide (eval.py:60-61), written by the loss
        kernel next to the hazards."""
        eng = self.last_engine
        return eng.buffer(0, "RISK")[:eng.B].clone()

    # ------------------------------------------------------------------------------------------------------------
    @torch.no_grad()
    def level_api(self, depth: int, data) -> Dict[str, torch.Tensor]:
        """RecursiveModel.forward(depth, PatchBatch) (model/interface.py:96-99 -> model/paths.py:66-146): packs the padded
        batch, runs one engine level, scatters the outputs back to the reference's padded layout."""
        dm = self.dims
        B, N = data.batch_size, data.max_patches
        D, Cc, d = dm.D, dm.C, dm.d
        num = data.num_ims.to(self.device)
        valid = data.valid_inds.to(self.device)
        M = int(num.sum().item())
        caps = [1] * dm.L
        maxseg = [1] * dm.L
        caps[depth] = M
        maxseg[depth] = N
        if depth > 0:
            caps[depth - 1] = M
            maxseg[depth - 1] = N
        eng = self.engine_for(B, caps, maxseg, [int(dm.K[i]) for i in range(dm.L - 1)], tag=f"level{depth}")
        seg = torch.zeros(B + 1, dtype=torch.int32, device=self.device)
        seg[1:] = torch.cumsum(num, 0).to(torch.int32)
        ar = torch.arange(B, dtype=torch.int32, device=self.device)
        eng.buffer(depth, "COUNT").fill_(M)
        eng.buffer(depth, "NTOK").fill_(M + B)
        eng.buffer(depth, "SEG").copy_(seg)
        eng.buffer(depth, "TSEG").copy_(seg[:B] + ar)
        row_slide = ar[:, None].expand(B, N)[valid]
        eng.buffer(depth, "ROW_SLIDE")[:M] = row_slide
        locs = torch.div(data.locs.to(self.device), self.model.model_config.patch_size, rounding_mode="floor")
        eng.buffer(depth, "ROW_XY").view(-1, 2)[:M] = locs[valid].to(torch.int32)
        eng.buffer(depth, "X").view(-1, D)[:M] = data.fts.to(self.device, torch.float32)[valid]
        lstm = bool(dm.lstm)
        if depth > 0:
            eng.buffer(depth, "ROW_PARENT")[:M] = torch.arange(M, dtype=torch.int32, device=self.device)
            if data.ctx_patch.shape[2] > 0:
                st = data.ctx_patch.to(self.device, torch.float32)[:, :, -1][valid]
                eng.buffer(depth - 1, "H").view(-1, D)[:M] = st[:, :D]
                if lstm:
                    eng.buffer(depth - 1, "CN").view(-1, Cc)[:M] = st[:, D:]
            cs = data.ctx_slide.to(self.device, torch.float32)
            if dm.residual == 2:                       # concat reads the whole slide-context history (paths.py:135)
                assert data.ctx_depth == depth, "slide_ctx_mode='concat' needs one slide context per previous level"
                for j in range(depth):
                    eng.buffer(j, "F").view(B, d).copy_(cs[:, j])
            elif data.ctx_depth > 0:
                eng.buffer(depth - 1, "F").view(B, d).copy_(cs[:, -1])
            else:
                eng.buffer(depth - 1, "F").zero_()
        else:
            eng.buffer(depth, "ROW_PARENT")[:M] = -1
        eng.refresh_weights()
        self._arm_dropout(eng)                          # model(depth, data) in train mode drops like the whole step does
        eng.level_forward(depth)
        out_imp = torch.zeros(B, N, device=self.device)
        out_imp[valid] = eng.buffer(depth, "ALPHA")[:M]
        ctx_patch = torch.zeros(B, N, D + Cc if lstm else D, device=self.device)
        if lstm:
            ctx_patch[valid] = torch.cat((eng.buffer(depth, "H").view(-1, D)[:M], eng.buffer(depth, "CN").view(-1, Cc)[:M]), dim=1)
        else:
            ctx_patch[valid] = eng.buffer(depth, "H").view(-1, D)[:M]
        return {"logits": eng.buffer(depth, "LOGITS").view(B, dm.nl).clone(),
                "ctx_slide": eng.buffer(depth, "F").view(B, d).clone(),
                "ctx_patch": ctx_patch, "importance": out_imp}


def _aggregate(self, depth: int, tokens: torch.Tensor, lengths: Optional[torch.Tensor]) -> torch.Tensor:
    """TransformerAggregator.forward of level `depth` on padded tokens [B, N, d] (model/aggregator.py:58-76)."""
    dm = self.dims
    B, N, d = tokens.shape
    assert d == dm.d
    dev = self.device
    num = (lengths if lengths is not None else torch.full((B,), N)).to(dev).long()
    valid = torch.arange(N, device=dev)[None, :] < num[:, None]
    M = int(num.sum().item())
    caps = [1] * dm.L; maxseg = [1] * dm.L
    caps[depth] = max(M, 1); maxseg[depth] = max(N, 1)
    eng = self.engine_for(B, caps, maxseg, [int(dm.K[i]) for i in range(dm.L - 1)], tag=f"agg{depth}")
    seg = torch.zeros(B + 1, dtype=torch.int32, device=dev)
    seg[1:] = torch.cumsum(num, 0).to(torch.int32)
    ar = torch.arange(B, dtype=torch.int32, device=dev)
    eng.buffer(depth, "COUNT").fill_(M); eng.buffer(depth, "NTOK").fill_(M + B)
    eng.buffer(depth, "SEG").copy_(seg); eng.buffer(depth, "TSEG").copy_(seg[:B] + ar)
    slide_of = ar[:, None].expand(B, N)[valid].long()
    rows = torch.arange(M, device=dev) + slide_of + 1                    # token row of patch row m of slide s: m + s + 1
    eng.buffer(depth, "TOK").view(-1, d)[rows] = tokens.detach().to(dev, torch.float32)[valid]
    eng.refresh_weights()
    self._arm_dropout(eng)
    L.check(eng.lib.paths_engine_aggregate(eng.handle, depth, torch.cuda.current_stream().cuda_stream), "engine_aggregate")
    return eng.buffer(depth, "Z").view(-1, d)[:B].clone()


class GraphStep:
    """One step of an engine captured as CUDA graphs: the forward pass (begin, L levels, selects, loss) and, optionally, the
    backward pass.  A step is ~270 dependent launches; at the reference's own batch size (32 slides) most of them run for a
    few microseconds and the host cannot enqueue them as fast as the GPU retires them -- replaying a captured graph removes
    the per-launch host cost.  The inputs live in static buffers (`slide_ids`, `ybin`, `cens`, and `fts` for host batches)
    that `load()` refills before a replay; every row count stays on the device, so the same graph serves any batch of the
    engine's size.  Dropout is baked into kernel parameters (the per-step seed), so graphs are used with dropout off."""

    def __init__(self, rt, eng, slide_ids, ybin, cens, scale, backward=True, fts_packed=None):
        self.rt, self.eng, self.scale, self.has_backward = rt, eng, scale, backward
        self.slide_ids, self.ybin, self.cens, self.fts = slide_ids.clone(), ybin.clone(), (cens.clone() if cens is not None else None), \
            (fts_packed.clone() if fts_packed is not None else None)
        eng.set_dropout(0.0, 0)
        st = torch.cuda.Stream(rt.device)
        st.wait_stream(torch.cuda.current_stream(rt.device))
        with torch.cuda.stream(st):                         # warm-up outside the capture: lazy kernel attributes, tensor maps
            for _ in range(2):
                self._enqueue_fwd()
                if backward:
                    eng.backward(True)
        torch.cuda.current_stream(rt.device).wait_stream(st)
        torch.cuda.synchronize(rt.device)
        l0 = eng.launch_count()
        self.g_fwd = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.g_fwd, stream=st, capture_error_mode="thread_local"):     # (a prefetch thread may be copying)
            self._enqueue_fwd()
        self.g_bwd = None
        if backward:
            self.g_bwd = torch.cuda.CUDAGraph()
            with torch.cuda.graph(self.g_bwd, stream=st, capture_error_mode="thread_local"):
                eng.backward(True)
        self.nodes = eng.launch_count() - l0
        rt.layout.attach_grads()

    def _enqueue_fwd(self):
        self.eng.forward(self.slide_ids, self.ybin, self.cens, self.scale, True, None, self.fts)

    def load(self, slide_ids=None, ybin=None, cens=None, fts_packed=None):
        if slide_ids is not None:
            self.slide_ids.copy_(slide_ids, non_blocking=True)
        if ybin is not None:
            self.ybin.copy_(ybin, non_blocking=True)
        if cens is not None and self.cens is not None:
            self.cens.copy_(cens, non_blocking=True)
        if fts_packed is not None:
            self.fts[:fts_packed.shape[0]].copy_(fts_packed, non_blocking=True)

    def forward(self):
        self.g_fwd.replay()
        return self.eng.overflow_poll()

    def backward(self):
        self.g_bwd.replay()

    def replay(self):
        """Whole step; raises on a capacity overflow (the caller re-runs eagerly on worst-case buffers)."""
        over, need = self.forward()
        if over:
            raise RuntimeError(f"paths_b200: graph step overflowed the level buffers ({need} rows needed)")
        if self.has_backward:
            self.backward()


class _NoCtx:
    """Stand-in for the autograd ctx when grad is disabled (torch.no_grad inference)."""
    def save_for_backward(self, *a):
        pass
