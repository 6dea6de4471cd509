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

    def _fetch_and_stage(self, it):
        return self._stage(next(it, None))

    def _stage(self, batch: Optional[Dict]):
        if batch is None:
            return None
        out = dict(batch)
        torch.cuda.set_device(self.device)                     # (helper thread: the current device is per thread)
        with torch.cuda.stream(self.copy_stream):
            # small (pageable) label tensors first: a pageable copy blocks the host until it has run, and behind the feature
            # copy that would be the whole 84 MB transfer
            have_packed = torch.is_tensor(batch.get("fts_packed"))
            for k in sorted(self.keys, key=lambda k: k in _BIG):
                v = batch.get(k)
                if k == "fts" and have_packed:
                    continue                                   # the step reads the packed rows: the padded copy stays on the host
                if torch.is_tensor(v) and not v.is_cuda:
                    if k in _BIG and not v.is_pinned():
                        v = v.pin_memory()                     # pageable memory would serialise the copy with the step
                    out[k] = v.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return out, ev

    def __iter__(self) -> Iterator[Dict]:
        it = iter(self.loader)
        pool = ThreadPoolExecutor(max_workers=1, thread_name_prefix="paths_b200_prefetch")
        try:
            pending = pool.submit(self._fetch_and_stage, it)
            while True:
                staged = pending.result()                      # (an exception of the loader or of a copy surfaces here)
                if staged is None:
                    break
                batch, ev = staged
                cur = torch.cuda.current_stream(self.device)
                cur.wait_event(ev)                             # the consumer's stream orders after the copy, no host sync
                for k in self.keys:
                    v = batch.get(k)
                    if torch.is_tensor(v) and v.is_cuda:
                        v.record_stream(cur)
                pending = pool.submit(self._fetch_and_stage, it)   # next batch: fetched and copied while this step runs
                yield batch
        finally:
            pool.shutdown(wait=True, cancel_futures=True)

    def __len__(self):
        return len(self.loader)
