"""Device-side input pipeline around the gather kernel.

Turns G1 from a kernel into the replacement for the reference's Dataset -> sampler -> collator -> ``.to(device)``
chain (FT.py:746, :757-772, :1030-1044 and HF Trainer's ``_prepare_inputs``): the permutation is generated on the host
with torch's CPU generator (bit-exact with the reference's ``random_split``), the epoch's index list is uploaded once,
and batches are produced by one gather launch each on a side stream, ``prefetch`` batches ahead of the consumer.
Honours ``--no-shuffle``, multiple epochs, the rank-strided distributed slice and a resume position.
"""
from __future__ import annotations

import collections
from typing import Iterator, Optional

import torch

from .dataset import TokenizedDataset
from .ordering import rank_slice, train_order


class DeviceBatcher:
    def __init__(self, ds: TokenizedDataset, batch_size: int, seed: int = 42, shuffle: bool = True, rank: int = 0,
                 world: int = 1, shift_labels: bool = True, prefetch: int = 2, drop_last: bool = False,
                 trainer_sampler: bool = False):
        if batch_size <= 0 or prefetch < 1:
            raise ValueError("batch_size and prefetch must be positive")
        self.ds, self.batch_size, self.seed, self.shuffle = ds, batch_size, seed, shuffle
        self.rank, self.world, self.shift_labels, self.prefetch, self.drop_last = rank, world, shift_labels, prefetch, drop_last
        self.stream = torch.cuda.Stream(ds.device)
        # FT.py:761: one permutation for the whole run (random_split happens once, before training)
        # trainer_sampler=True additionally re-draws the per-epoch DistributedSampler permutation the HF Trainer
        # applies on top of it (ordering.sampler_epoch_order); the default walks the split order itself.
        self.trainer_sampler = trainer_sampler
        self._split = train_order(len(ds), seed, shuffle)
        self.set_epoch(0)

    def set_epoch(self, epoch: int) -> None:
        if self.trainer_sampler:
            from .ordering import sampler_epoch_order
            order = sampler_epoch_order(self._split, self.seed, epoch, self.rank, self.world, shuffle=True)
        else:
            order = rank_slice(self._split, self.rank, self.world)
        self.order_host = order.contiguous()
        self.order_dev = self.order_host.to(self.ds.device)  # 8 B per context

    def batches_per_epoch(self) -> int:
        n = self.order_host.numel()
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    def index_list(self, batch_no: int) -> torch.Tensor:
        """Host-side view of which contexts form batch ``batch_no`` of an epoch (for logging / resume checks)."""
        i = batch_no * self.batch_size
        return self.order_host[i:i + self.batch_size]

    def _launch(self, batch_no: int):
        i = batch_no * self.batch_size
        idx = self.order_dev[i:i + self.batch_size]
        with torch.cuda.stream(self.stream):
            b = self.ds.batch(idx, self.shift_labels)
            ev = torch.cuda.Event()
            ev.record(self.stream)
        return b, ev

    def epoch(self, start_batch: int = 0) -> Iterator[dict]:
        """Yields this rank's batches of one epoch, starting at ``start_batch`` (resume position).  Each batch is
        ready on the CURRENT stream when it is yielded; the next ``prefetch`` batches are already in flight."""
        n = self.batches_per_epoch()
        queue = collections.deque()
        nxt = start_batch
        cur = torch.cuda.current_stream(self.ds.device)
        self.stream.wait_stream(cur)  # the token stream upload / previous consumers are ordered before the gathers
        while nxt < n and len(queue) < self.prefetch:
            queue.append(self._launch(nxt))
            nxt += 1
        while queue:
            b, ev = queue.popleft()
            if nxt < n:
                queue.append(self._launch(nxt))
                nxt += 1
            cur.wait_event(ev)
            for t in b.values():
                t.record_stream(cur)  # allocated on the side stream, consumed on the current one
            yield b
