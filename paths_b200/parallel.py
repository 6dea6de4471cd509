"""Slide-level data parallelism for the PATHS step (new work -- the reference is single-process).

Slides never interact in the forward or backward pass, so a global batch is split by slide across ranks (one process per
GPU, each with a full parameter replica and its own shard of the pyramid store) and the only exchange is one
all-reduce(sum) of the flat fp32 gradient buffer per step.  The buffer is cut into buckets in BACKWARD order --
procs[L-1], ..., procs[0], then the shared LSTM (whose gradient is complete only after level 0) -- and each bucket's
all-reduce is enqueued behind the event the engine records when that level's gradients are final, so NCCL (its own stream,
NVLink / NVSwitch) overlaps the remaining backward.  The loss is a mean over the GLOBAL batch: every rank scales by
1 / global_batch (unequal shards stay exact), so the reduction is a plain sum.
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.distributed as dist

from . import _lib as L


class GradBuckets:
    """Contiguous [start, end) ranges (in floats) of the flat gradient buffer, in the order their gradients complete."""

    def __init__(self, offsets: Sequence[int], total: int, num_levels: int, has_lstm: bool = True):
        first = [int(offsets[L.P_LSTM_COUNT + l * L.LEVEL_STRIDE + L.P_WUV]) for l in range(num_levels)]
        assert all(f >= 0 for f in first) and first == sorted(first), "levels must be laid out in ascending order"
        ends = first[1:] + [int(total)]
        self.ranges: List[Tuple[str, int, int]] = []
        for l in reversed(range(num_levels)):
            self.ranges.append((f"procs.{l}", first[l], ends[l]))
        if has_lstm and first[0] > 0:
            self.ranges.append(("lstm", 0, first[0]))
        else:
            assert first[0] == 0, "flat buffer must start with the LSTM group or with level 0"
        self.total = int(total)

    @staticmethod
    def from_layout(layout) -> "GradBuckets":
        return GradBuckets(list(layout.offsets), layout.total, layout.dims.L, bool(layout.dims.lstm))

    def covers_exactly_once(self) -> bool:
        spans = sorted((a, b) for _, a, b in self.ranges)
        pos = 0
        for a, b in spans:
            if a != pos or b <= a:
                return False
            pos = b
        return pos == self.total

    def level_bucket(self, level: int) -> Tuple[int, int]:
        for name, a, b in self.ranges:
            if name == f"procs.{level}":
                return a, b
        raise KeyError(level)

    def lstm_bucket(self) -> Optional[Tuple[int, int]]:
        for name, a, b in self.ranges:
            if name == "lstm":
                return a, b
        return None

    def reduce_all(self, gflat: torch.Tensor, group=None, async_op: bool = False):
        """All-reduce every bucket in backward order (used after a monolithic backward, and by the CPU tests)."""
        works = [dist.all_reduce(gflat[a:b], op=dist.ReduceOp.SUM, group=group, async_op=True) for _, a, b in self.ranges]
        if async_op:
            return works
        for w in works:
            w.wait()
        return []


class DataParallelStep:
    """One training step across ranks: forward, whole-step backward, bucketed all-reduce overlapped with it.

        dp = DataParallelStep(model)                      # after dist.init_process_group("nccl")
        hazards, loss = dp.step(batch, top_k)             # batch = this rank's shard (collated, "slide" entries)
        opt.step()

    or, keeping the reference's loop body (train.py:58-66) unchanged:

        dp = DataParallelStep(model).attach(global_batch=B * world)
        hazards, loss = utils.inference_end2end(...); loss.backward(); opt.step()

    The backward pass is the engine's whole-step one (grouped aggregators on their own streams); it records an event when
    the gradients of procs[l] are final, and each bucket's all-reduce is enqueued behind its event on a side stream, so NCCL
    runs while the shallower levels are still computing.  `loss` is this rank's share of the global mean loss (sum over ranks
    = global loss); `dp.global_loss(loss)` reduces it.  Gradients land, summed over ranks, in the parameters' .grad (views of
    the flat buffer).  `last_exposed_comm()` returns the device time between the end of the backward pass and the end of the
    last all-reduce of the latest step: the communication that was NOT hidden.
    """

    def __init__(self, model, group=None, overlap: bool = True):
        self.model = model
        self.group = group
        self.world = dist.get_world_size(group) if dist.is_initialized() else 1
        self.rt = model.runtime()
        self.buckets = GradBuckets.from_layout(self.rt.layout)
        assert self.buckets.covers_exactly_once()
        self.overlap = overlap
        self.global_batch: Optional[int] = None
        self._side = torch.cuda.Stream(self.rt.device) if torch.cuda.is_available() else None
        self._ev = None

    def attach(self, global_batch: Optional[int] = None) -> "DataParallelStep":
        """Route `loss.backward()` of utils.inference_end2end through this object's backward (all-reduce included) and scale the
        loss by 1 / global_batch (None: local batch x world size)."""
        self.global_batch = global_batch
        self.rt.grad_sync = self
        return self

    def detach(self) -> None:
        if self.rt.grad_sync is self:
            self.rt.grad_sync = None

    def step(self, batch: Dict, keep_patches, global_batch: Optional[int] = None):
        rt = self.rt
        B = len(batch["slide"])
        if global_batch is None:
            if self.world > 1:
                t = torch.tensor([B], device=rt.device, dtype=torch.int64)
                dist.all_reduce(t, group=self.group)
                global_batch = int(t.item())
            else:
                global_batch = B
        keep = [keep_patches] * (rt.dims.L - 1) if isinstance(keep_patches, int) else list(keep_patches)
        with torch.no_grad():
            hz, loss = rt.end2end(batch, keep, global_batch=global_batch)
        eng = rt.last_engine
        self.backward(eng)
        return hz, loss

    def backward(self, eng, zero: bool = True) -> None:
        """loss.backward() + gradient all-reduce.  zero=False: the flat gradient buffer was prepared by the caller (autograd
        accumulation semantics)."""
        rt = self.rt
        gflat = rt.layout.gflat
        nl = eng.dims.L
        eng.backward(zero)                                  # whole step, enqueued at once; ev_grad[l] recorded inside
        if self.world > 1:
            cur = torch.cuda.current_stream(rt.device)
            timing = self._ev is not None
            if timing:
                self._ev[0].record(cur)                     # end of this rank's backward compute
            works = []
            if self.overlap:
                with torch.cuda.stream(self._side):
                    for l in reversed(range(nl)):
                        eng.wait_grads(l, self._side)       # NCCL orders after the side stream, i.e. after the bucket's event
                        a, b = self.buckets.level_bucket(l)
                        works.append(dist.all_reduce(gflat[a:b], group=self.group, async_op=True))
                    lb = self.buckets.lstm_bucket()
                    if lb is not None:                      # the shared LSTM's gradient is final after level 0
                        works.append(dist.all_reduce(gflat[lb[0]:lb[1]], group=self.group, async_op=True))
            else:
                works.append(dist.all_reduce(gflat, group=self.group, async_op=True))
            for w in works:
                w.wait()                                    # the current stream waits for NCCL's; the host does not block
            if timing:
                self._ev[1].record(cur)
        rt.layout.attach_grads()

    def time_exposed_comm(self, on: bool = True) -> None:
        self._ev = (torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) if on else None

    def last_exposed_comm(self) -> float:
        """Milliseconds between the end of the latest backward pass and the end of its last all-reduce (synchronises)."""
        if self._ev is None or self.world == 1:
            return 0.0
        self._ev[1].synchronize()
        return self._ev[0].elapsed_time(self._ev[1])

    def global_loss(self, loss: torch.Tensor) -> torch.Tensor:
        out = loss.detach().clone()
        if self.world > 1:
            dist.all_reduce(out, group=self.group)
        return out


def shard_slides(n_slides: int, rank: int, world: int, weights: Optional[Sequence[int]] = None) -> List[int]:
    """Slide indices owned by `rank`.  Without weights: contiguous equal shards.  With weights (level-0 bag sizes, the only
    unbounded term): greedy longest-first balancing so a 20 k-cell slide does not serialise one rank."""
    if weights is None:
        per = (n_slides + world - 1) // world
        return list(range(rank * per, min(n_slides, (rank + 1) * per)))
    order = sorted(range(n_slides), key=lambda i: -int(weights[i]))
    load = [0] * world
    owner = [0] * n_slides
    for i in order:
        r = min(range(world), key=lambda k: (load[k], k))
        owner[i] = r
        load[r] += int(weights[i])
    return [i for i in range(n_slides) if owner[i] == rank]
