"""Host -> device staging one batch ahead of the step that consumes it.

The reference moves each collated batch to the GPU inside the step (`PatchBatch.from_batch(batch, device)`,
utils.py:236; train.py:58-60 iterates the DataLoader directly), so the level-0 feature copy (84 MB for 256 BRCA-shaped
slides) sits on the critical path of every step.  `DevicePrefetcher` wraps the same iterable of collated batches and
issues the copy of batch i+1 on a dedicated copy stream while step i computes; `utils.inference_end2end` accepts the
yielded batch unchanged (its `.to(device)` calls are no-ops on tensors that already live there).  Fetching the next batch
from the loader and enqueuing its copies happens on a helper thread: when the consumer comes back from reading its loss the
GPU is idle, and every host millisecond before the next step's first launch is a millisecond of idle GPU.

    for batch in DevicePrefetcher(train_loader, device):       # instead of: for batch in train_loader
        hazards, loss = utils.inference_end2end(..., batch, config.task)
"""
from __future__ import annotations

from concurrent.futures import ThreadPoolExecutor
from typing import Dict, Iterable, Iterator, Optional

import torch

_TENSOR_KEYS = ("fts", "fts_packed", "survival_bin", "censored", "subtype")
_BIG = ("fts", "fts_packed")


class DevicePrefetcher:
    def __init__(self, loader: Iterable[Dict], device, keys=_TENSOR_KEYS):
        self.loader = loader
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("paths_b200: DevicePrefetcher stages batches into CUDA memory; there is no CPU path")
        if self.device.index is None:
            self.device = torch.device("cuda", torch.cuda.current_device())
        self.keys = tuple(keys)
        self.copy_stream = torch.cuda.Stream(self.device)
        # The big tensors (level-0 features) land in a ring of RING device buffers owned by the prefetcher instead of a fresh
        # allocation per step: no allocator traffic (and no cudaMalloc stall) on the step path.  Slot k is refilled only after the
        # consumer has come back for the batch after next -- everything that read it has been enqueued by then -- and the copy
        # stream waits on the event recorded at that moment.
        self._ring = {}                  # key -> list of RING flat buffers
        self._done = [None] * self.RING  # event recorded on the consumer's stream when it has finished enqueuing slot k's step
        self._count = 0

    RING = 3

    def _slot_buffer(self, key, slot, like: torch.Tensor) -> torch.Tensor:
        bufs = self._ring.setdefault(key, [None] * self.RING)
        b = bufs[slot]
        if b is None or b.numel() < like.numel() or b.dtype != like.dtype:
            b = bufs[slot] = torch.empty(like.numel(), dtype=like.dtype, device=self.device)
        return b[:like.numel()].view(like.shape)

    def _fetch_and_stage(self, it, slot):
        return self._stage(next(it, None), slot)

    def _stage(self, batch: Optional[Dict], slot: int = 0):
        if batch is None:
            return None
        out = dict(batch)
        torch.cuda.set_device(self.device)                     # (helper thread: the current device is per thread)
        with torch.cuda.stream(self.copy_stream):
            if self._done[slot] is not None:
                self.copy_stream.wait_event(self._done[slot])  # the slot's previous consumer
            # small (pageable) label tensors first: a pageable copy blocks the host until it has run, and behind the feature
            # copy that would be the whole 84 MB transfer
            have_packed = torch.is_tensor(batch.get("fts_packed"))
            for k in sorted(self.keys, key=lambda k: k in _BIG):
                v = batch.get(k)
                if k == "fts" and have_packed:
                    continue                                   # the step reads the packed rows: the padded copy stays on the host
                if torch.is_tensor(v) and not v.is_cuda:
                    if k in _BIG:
                        if not v.is_pinned():
                            v = v.pin_memory()                 # pageable memory would serialise the copy with the step
                        dst = self._slot_buffer(k, slot, v)
                        dst.copy_(v, non_blocking=True)
                        out[k] = dst
                    else:
                        out[k] = v.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return out, ev

    def __iter__(self) -> Iterator[Dict]:
        it = iter(self.loader)
        pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="paths_b200_prefetch")
        try:
            slot = 0
            pending = pool.submit(self._fetch_and_stage, it, slot)
            while True:
                staged = pending.result()                      # (an exception of the loader or of a copy surfaces here)
                if staged is None:
                    break
                batch, ev = staged
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)                             # the consumer's stream orders after the copy, no host sync
                for k in self.keys:
                    v = batch.get(k)
                    if torch.is_tensor(v) and v.is_cuda and k not in _BIG:
                        v.record_stream(cur)
                used, slot = slot, (slot + 1) % self.RING
                pending = pool.submit(self._fetch_and_stage, it, slot)   # next batch: fetched and copied while this step runs
                yield batch
                done = torch.cuda.Event()                      # the consumer is back: this batch's step has been enqueued
                done.record(torch.cuda.current_stream(self.device))
                self._done[used] = done
        finally:
            pool.shutdown(wait=True, cancel_futures=True)

    def __len__(self):
        return len(self.loader)
