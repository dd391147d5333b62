"""Device-resident mirror of the reference's dataset/collate surface.

Reference: ``TokenizedDataset`` FT.py:633-695 (mmap of a flat uint16 token file, ``load`` slices one context, widens
to int64 and builds the bool mask) and ``collector`` FT.py:1030-1035 (three stacks).  Here the token stream is
uploaded to HBM once and a batch is built by ONE kernel launch (ftk_gather_windows) from an index list.
Same names, argument meaning and results; tensors come back on the GPU.
"""
from __future__ import annotations

import os
from typing import List, Optional, Sequence, Tuple, Union

import numpy as np
import torch

from . import ops


class TokenizedDataset(torch.utils.data.Dataset):
    """``TokenizedDataset(path, context_length=2048)`` as in the reference; the tokenizer's pad/eos ids, which the
    reference reads from a module global (FT.py:650-651), are passed explicitly (``tokenizer=`` or the two ids)."""

    def __init__(self, path: str, context_length: int = 2048, tokenizer=None, pad_token_id: Optional[int] = None,
                 eos_token_id: Optional[int] = None, device: Union[str, torch.device, None] = None):
        if context_length <= 0:
            raise ValueError("context_length must be positive")
        if tokenizer is not None:
            pad_token_id, eos_token_id = tokenizer.pad_token_id, tokenizer.eos_token_id
        if pad_token_id is None or eos_token_id is None:
            raise ValueError("pass tokenizer= or pad_token_id= and eos_token_id=")
        file_stat = os.stat(path)  # raises like the reference when the file is missing
        self._context_length = context_length
        self.length = int(file_stat.st_size / 2 / context_length)  # FT.py:647: trailing partial context dropped
        self._padding_is_ambiguous = pad_token_id == eos_token_id  # FT.py:650
        self._pad_token_id = int(pad_token_id)
        self.device = torch.device(device if device is not None else f"cuda:{torch.cuda.current_device()}")
        n_tok = self.length * context_length
        host = np.fromfile(path, dtype=np.uint16, count=n_tok)  # native order, as numpy.ushort at FT.py:669
        pinned = torch.from_numpy(host.view(np.int16))
        self.tokens = pinned.to(self.device, non_blocking=False)  # int16 view of the uint16 stream, 2 B/token

    def __len__(self) -> int:
        return self.length

    def _index_tensor(self, idx: Union[int, Sequence[int], torch.Tensor]) -> torch.Tensor:
        t = torch.as_tensor(idx, dtype=torch.int64).reshape(-1)
        if t.device.type == "cpu":
            if t.numel() and (int(t.min()) < 0 or int(t.max()) >= self.length):
                raise IndexError(f"context index out of range [0, {self.length})")
            t = t.to(self.device, non_blocking=True)
        return t.contiguous()

    def batch(self, idx, shift_labels: bool = False) -> dict:
        """One launch: {"input_ids", "attention_mask", "labels", "n_valid"} for the contexts in ``idx``.
        ``labels`` already carry -100 where the mask is False (FT.py:476-477); ``shift_labels=True`` also applies
        HF's causal shift so they can feed the fused loss directly."""
        ids, labels, mask, n_valid = torch.ops.ftk.gather_windows(
            self.tokens, self._index_tensor(idx), self._context_length, self._pad_token_id,
            self._padding_is_ambiguous, shift_labels)
        return {"input_ids": ids, "attention_mask": mask, "labels": labels, "n_valid": n_valid}

    def load(self, idx: int) -> Tuple[torch.Tensor, torch.Tensor]:
        b = self.batch([int(idx)])
        return b["input_ids"][0], b["attention_mask"][0]

    def __getitem__(self, idx) -> Tuple[torch.Tensor, torch.Tensor]:
        return self.load(idx)

    def index_view(self) -> "IndexView":
        """A dataset whose items are bare context indices, for use with ``DeviceCollator`` under HF Trainer."""
        return IndexView(self)


class IndexView(torch.utils.data.Dataset):
    def __init__(self, ds: TokenizedDataset):
        self.ds = ds

    def __len__(self) -> int:
        return len(self.ds)

    def __getitem__(self, idx: int) -> int:
        if not (0 <= idx < len(self.ds)):
            raise IndexError(idx)
        return int(idx)


def collector(data: List[Tuple[torch.Tensor, torch.Tensor]]) -> dict:
    """FT.py:1030-1035 for items already materialised by ``TokenizedDataset.__getitem__`` (device tensors)."""
    return {
        "input_ids": torch.stack([f[0] for f in data]),
        "attention_mask": torch.stack([f[1] for f in data]),
        "labels": torch.stack([f[0] for f in data]),
    }


class DeviceCollator:
    """``data_collator`` for ``IndexView`` items: the whole batch in one gather launch."""

    def __init__(self, ds: TokenizedDataset, shift_labels: bool = False):
        self.ds, self.shift_labels = ds, shift_labels

    def __call__(self, indices: List[int]) -> dict:
        b = self.ds.batch(indices, self.shift_labels)
        b.pop("n_valid")
        return b
