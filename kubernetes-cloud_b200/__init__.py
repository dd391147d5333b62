"""B200-native (sm_100a) training hot path of kubernetes-cloud's finetuner-workflow/finetuner.

Drop-in pieces, named after the reference surfaces they replace (finetuner-workflow/finetuner/finetuner.py):
  TokenizedDataset, collector   -> dataset.py   (FT.py:633-695, :1030-1035)
  compute_loss / ModifiedTrainerMixin -> loss.py (FT.py:469-493)
  train_order, rank_slice       -> ordering.py  (FT.py:739, :757-772 + the Trainer's distributed sampler)
  GradSync                      -> dp.py        (ds_config.json:27-45 gradient reduction)
All compute goes through lib/libftk.so (C ABI in include/ftk.h); nothing here falls back to CPU or eager code.
"""
from . import _lib  # noqa: F401
from ._lib import FtkError, build  # noqa: F401

__all__ = ["FtkError", "build", "TokenizedDataset", "collector", "DeviceCollator", "compute_loss",
           "ModifiedTrainerMixin", "fused_lmhead_ce", "train_order", "rank_slice", "GradSync", "FusedAdamW", "WarmupLR",
           "DeviceBatcher", "write_token_file", "validate_token_file"]


def __getattr__(name):  # lazy: importing the package must not need torch.cuda or the .so
    if name in ("TokenizedDataset", "collector", "DeviceCollator"):
        from . import dataset
        return getattr(dataset, name)
    if name in ("compute_loss", "ModifiedTrainerMixin", "shift_labels"):
        from . import loss
        return getattr(loss, name)
    if name == "fused_lmhead_ce":
        from . import ops
        return ops.fused_lmhead_ce
    if name in ("train_order", "rank_slice", "epoch_batches"):
        from . import ordering
        return getattr(ordering, name)
    if name in ("GradSync",):
        from . import dp
        return getattr(dp, name)
    if name in ("FusedAdamW", "WarmupLR"):
        from . import optim
        return getattr(optim, name)
    if name in ("DeviceBatcher",):
        from . import pipeline
        return getattr(pipeline, name)
    if name in ("write_token_file", "validate_token_file", "TokenFileReport"):
        from . import tokenfile
        return getattr(tokenfile, name)
    raise AttributeError(name)
