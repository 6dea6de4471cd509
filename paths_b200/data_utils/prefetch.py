"""Host -> device staging one batch ahead of the step that consumes it.

The reference moves each collated batch to the GPU inside the step (`PatchBatch.from_batch(batch, device)`,
utils.py:236; train.py:58-60 iterates the DataLoader directly), so the level-0 feature copy (84 MB for 256 BRCA-shaped
slides) sits on the critical path of every step.  `DevicePrefetcher` wraps the same iterable of collated batches and
issues the copy of batch i+1 on a dedicated copy stream while step i computes; `utils.inference_end2end` accepts the
yielded batch unchanged (its `.to(device)` calls are no-ops on tensors that already live there).

    for batch in DevicePrefetcher(train_loader, device):       # instead of: for batch in train_loader
        hazards, loss = utils.inference_end2end(..., batch, config.task)
"""
from __future__ import annotations

from typing import Dict, Iterable, Iterator, Optional

import torch

_TENSOR_KEYS = ("fts", "survival_bin", "censored", "subtype")


class DevicePrefetcher:
    def __init__(self, loader: Iterable[Dict], device, keys=_TENSOR_KEYS):
        self.loader = loader
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("paths_b200: DevicePrefetcher stages batches into CUDA memory; there is no CPU path")
        self.keys = tuple(keys)
        self.copy_stream = torch.cuda.Stream(self.device)

    def _stage(self, batch: Optional[Dict]):
        if batch is None:
            return None
        out = dict(batch)
        with torch.cuda.stream(self.copy_stream):
            # small (pageable) label tensors first: a pageable copy blocks the host until it has run, and behind the feature
            # copy that would be the whole 84 MB transfer
            for k in sorted(self.keys, key=lambda k: k == "fts"):
                v = batch.get(k)
                if torch.is_tensor(v) and not v.is_cuda:
                    if k == "fts" and not v.is_pinned():
                        v = v.pin_memory()                     # pageable memory would serialise the copy with the step
                    out[k] = v.to(self.device, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self.copy_stream)
        return out, ev

    def __iter__(self) -> Iterator[Dict]:
        it = iter(self.loader)
        staged = self._stage(next(it, None))
        while staged is not None:
            batch, ev = staged
            cur = torch.cuda.current_stream(self.device)
            cur.wait_event(ev)                                 # the consumer's stream orders after the copy, no host sync
            for k in self.keys:
                v = batch.get(k)
                if torch.is_tensor(v) and v.is_cuda:
                    v.record_stream(cur)
            staged = self._stage(next(it, None))               # copy of the next batch overlaps this step's kernels
            yield batch

    def __len__(self):
        return len(self.loader)
