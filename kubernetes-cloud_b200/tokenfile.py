"""The token-file format contract, host side.

The format is defined only by the reference's reader: a flat file of native-endian ``uint16`` token ids, aligned
along ``context_length`` chunks (``TokenizedDataset`` docstring FT.py:634-637; ``length = int(size / 2 / T)`` FT.py:647
silently drops a trailing partial context), written by an external tokenizer the workflow runs with ``-context``,
``-pad`` and ``-eot`` (finetune-workflow.yaml:441-454).  The mask rule depends on a convention of that writer: when the
pad token equals the end-of-text token, only the FINAL context may be padded (FT.py:674-687).  ``write_token_file``
produces files that honour the contract; ``validate_token_file`` checks a file against it so synthetic and real data
are interchangeable.
"""
from __future__ import annotations

import dataclasses
import os
from typing import List, Optional, Sequence

import numpy as np


@dataclasses.dataclass
class TokenFileReport:
    path: str
    size_bytes: int
    context_length: int
    n_contexts: int
    trailing_bytes: int            # bytes after the last whole context (the reader drops them)
    max_token: int
    vocab_ok: Optional[bool]       # None when no vocab size was given
    padded_contexts: List[int]     # contexts that end in >= 2 pad tokens (only meaningful with pad_id)
    convention_ok: Optional[bool]  # pad == eos: only the final context may be padded
    problems: List[str]

    @property
    def ok(self) -> bool:
        return not self.problems


def write_token_file(path: str, tokens: Sequence[int], context_length: int, pad_id: int) -> int:
    """Writes ``tokens`` as uint16 and pads the final context with ``pad_id`` to a whole number of contexts
    (what the reference's tokenizer does).  Returns the number of contexts written."""
    arr = np.asarray(tokens)
    if arr.size and (arr.min() < 0 or arr.max() > 0xFFFF):
        raise ValueError("token ids must fit in uint16 (FT.py:669 reads numpy.ushort)")
    if not (0 <= pad_id <= 0xFFFF):
        raise ValueError("pad_id must fit in uint16")
    if context_length <= 0:
        raise ValueError("context_length must be positive")
    arr = arr.astype(np.uint16)
    rem = arr.size % context_length
    if rem:
        arr = np.concatenate([arr, np.full(context_length - rem, pad_id, dtype=np.uint16)])
    arr.tofile(path)
    return arr.size // context_length


def validate_token_file(path: str, context_length: int, vocab_size: Optional[int] = None, pad_id: Optional[int] = None,
                        eos_id: Optional[int] = None, chunk_contexts: int = 4096) -> TokenFileReport:
    size = os.stat(path).st_size
    n_ctx = size // (2 * context_length)
    trailing = size - n_ctx * 2 * context_length
    problems: List[str] = []
    if size % 2:
        problems.append("file size is odd: not a uint16 stream")
    if trailing:
        problems.append(f"{trailing} trailing bytes do not fill a context of {context_length} tokens "
                        "(the reader drops them, FT.py:647)")
    if n_ctx == 0:
        problems.append("file holds no whole context")
    max_tok = -1
    padded: List[int] = []
    if n_ctx:
        mm = np.memmap(path, dtype=np.uint16, mode="r", shape=(n_ctx, context_length))
        for c0 in range(0, n_ctx, chunk_contexts):
            blk = np.asarray(mm[c0:c0 + chunk_contexts])
            max_tok = max(max_tok, int(blk.max()))
            if pad_id is not None and context_length >= 2:
                tail = (blk[:, -1] == pad_id) & (blk[:, -2] == pad_id)
                padded += [int(c0 + i) for i in np.nonzero(tail)[0]]
        del mm
    vocab_ok = None
    if vocab_size is not None:
        vocab_ok = max_tok < vocab_size
        if not vocab_ok:
            problems.append(f"token id {max_tok} is outside the vocabulary of {vocab_size}")
    convention_ok = None
    if pad_id is not None and eos_id is not None and pad_id == eos_id:
        convention_ok = all(c == n_ctx - 1 for c in padded)
        if not convention_ok:
            bad = [c for c in padded if c != n_ctx - 1][:8]
            problems.append(f"pad == eos but non-final contexts {bad} end in padding: their pads would be trained on "
                            "(FT.py:674-687 masks only the final context)")
    return TokenFileReport(path, size, context_length, n_ctx, trailing, max_tok, vocab_ok, padded, convention_ok, problems)


def synthetic_token_stream(n_contexts: int, context_length: int, vocab_size: int, seed: int, pad_id: int = 50256,
                           trailing_pads: int = 300) -> np.ndarray:
    """The synthetic stream SURVEY.md 8(d) prescribes for benchmarks: uniform uint16 ids from a seeded numpy generator,
    the final context ending in ``trailing_pads`` pad tokens so the mask rule is exercised."""
    rng = np.random.default_rng(seed)
    tok = rng.integers(0, vocab_size, size=n_contexts * context_length, dtype=np.uint16)
    if trailing_pads:
        tok[n_contexts * context_length - min(trailing_pads, context_length):] = pad_id
    return tok
