"""Which contexts form which batch -- host side, integer exact.

Mirrors FT.py:739 (torch.manual_seed(seed)), FT.py:757-772 (random_split permutation, or file order with
--no-shuffle; train_ratio is forced to 1 at FT.py:748-756) and the rank-strided slice the HF Trainer's
DistributedSampler applies for world > 1 (torch/utils/data/distributed.py:107-134).  The permutation is drawn
with torch's CPU generator, as the reference does; only the index list (8 B per context) goes to the device.
"""
from __future__ import annotations

from typing import Iterator

import torch


def train_order(n_ctx: int, seed: int = 42, shuffle: bool = True) -> torch.Tensor:
    """int64 [n_ctx]: the order in which the reference's train_dataset enumerates contexts."""
    if n_ctx < 0:
        raise ValueError("n_ctx must be >= 0")
    if not shuffle:
        return torch.arange(n_ctx, dtype=torch.int64)
    g = torch.Generator()
    g.manual_seed(seed)
    return torch.randperm(n_ctx, generator=g)


def rank_slice(order: torch.Tensor, rank: int, world: int) -> torch.Tensor:
    """order[rank::world] after padding by wrap-around to a multiple of world (DistributedSampler, drop_last=False)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} not in [0, {world})")
    n = order.numel()
    if n == 0 or world == 1:
        return order
    total = (n + world - 1) // world * world
    pad = total - n
    if pad:
        reps = (pad + n - 1) // n
        order = torch.cat([order, order.repeat(reps)[:pad]])
    return order[rank:total:world]


def epoch_batches(n_ctx: int, batch_size: int, seed: int = 42, shuffle: bool = True, rank: int = 0, world: int = 1,
                  drop_last: bool = False) -> Iterator[torch.Tensor]:
    """Yields this rank's int64 index lists, one per micro-batch."""
    mine = rank_slice(train_order(n_ctx, seed, shuffle), rank, world)
    for i in range(0, mine.numel(), batch_size):
        b = mine[i:i + batch_size]
        if drop_last and b.numel() < batch_size:
            return
        yield b


def sampler_epoch_order(train_indices: torch.Tensor, seed: int, epoch: int, rank: int, world: int,
                        shuffle: bool = True) -> torch.Tensor:
    """The order in which HF Trainer's ``DistributedSampler`` (world > 1, ``seed=args.seed``) walks the reference's
    ``train_dataset`` in a given epoch, composed with the ``random_split`` permutation behind that Subset:
    positions = randperm(n) from a generator seeded ``seed + epoch`` (or 0..n-1 without shuffle), padded by wrap-around
    to a multiple of ``world``, sliced ``rank::world`` (torch/utils/data/distributed.py:107-134), then mapped through
    ``train_indices`` (FT.py:761).  For world == 1 the Trainer uses a RandomSampler seeded from the global RNG state
    at that moment, which depends on how many draws model initialisation made and cannot be reproduced from the
    flags alone; this function applies the same seeded rule there."""
    n = train_indices.numel()
    if shuffle:
        g = torch.Generator()
        g.manual_seed(seed + epoch)
        pos = torch.randperm(n, generator=g)
    else:
        pos = torch.arange(n)
    return train_indices[rank_slice(pos, rank, world)]
