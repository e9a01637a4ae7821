"""Dataset drop-ins with the reference's surface: ``Split`` (``torchbooster/dataset.py:15-22``) and
``BaseDataset`` (``:25-73``), backed by this package's own read-only LMDB reader."""
from __future__ import annotations

from enum import Enum
from pathlib import Path
from typing import Any

from torch.nn import Module
from torch.utils.data import Dataset

from .lmdb import LMDBReader

Split = Enum("Split", {"TRAIN": "train", "VALID": "validation", "TEST": "test"})
Split.__doc__ = "Dataset split: ``TRAIN`` / ``VALID`` / ``TEST`` (values 'train', 'validation', 'test')."


class BaseDataset(Dataset):
    """Root of LMDB-backed datasets.

    The constructor only records its arguments and builds ``self.lmdb_reader`` (which opens the
    database lazily, once per process).  A concrete dataset decodes ``self.lmdb_reader[idx]`` in
    ``__getitem__`` and applies ``self.transform``; ``prepare`` builds the database.  Both raise
    ``NotImplementedError`` here.  ``len(dataset)`` is the database's ``length`` record.

    ``map_size`` keeps the reference default of ``1024 * 4`` (``torchbooster/dataset.py:49``); it
    is irrelevant for a read-only mapping, as is ``max_readers``.
    """

    def __init__(self, path: Path, transform: Module = None, map_size: int = 1024 * 4, max_readers: int = 126) -> None:
        super().__init__()
        self.path, self.transform = path, transform
        self.map_size, self.max_readers = map_size, max_readers
        self.lmdb_reader = LMDBReader(path, map_size=map_size, max_readers=max_readers)

    def __len__(self) -> int:
        return len(self.lmdb_reader)

    def __getitem__(self, idx: int) -> Any:
        raise NotImplementedError("Method '__getitem__' is not implemented.")

    @classmethod
    def prepare(cls, *args, **kwargs) -> None:
        raise NotImplementedError("Method 'prepare' is not implemented.")


__all__ = ["BaseDataset", "Split"]
