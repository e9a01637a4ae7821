"""Dataset drop-ins: ``Split`` and ``BaseDataset`` with the signatures of
``torchbooster/dataset.py:15-22`` and ``:25-73``."""
from __future__ import annotations

from enum import Enum
from pathlib import Path
from typing import Any

from torch.nn import Module
from torch.utils.data import Dataset

from .lmdb import LMDBReader


class Split(Enum):
    """Dataset split (train, validation, test)."""
    TRAIN = "train"
    VALID = "validation"
    TEST = "test"


class BaseDataset(Dataset):
    """Base class of LMDB-backed datasets.

    Holds an ``LMDBReader`` over ``path`` (opened lazily, per process) and a ``transform``;
    subclasses implement ``__getitem__`` (decode ``self.lmdb_reader[idx]``, apply
    ``self.transform``) and ``prepare``.  Defaults follow the reference, including its
    ``map_size = 1024 * 4`` (``torchbooster/dataset.py:49``), which is harmless for a read-only map.
    """

    def __init__(
        self,
        path: Path,
        transform: Module = None,
        map_size: int = 1024 * 4,
        max_readers: int = 126,
    ) -> None:
        super().__init__()
        self.path = path
        self.transform = transform
        self.map_size = map_size
        self.max_readers = max_readers
        self.lmdb_reader = LMDBReader(self.path, map_size=self.map_size, max_readers=self.max_readers)

    def __len__(self) -> int:
        return len(self.lmdb_reader)

    def __getitem__(self, idx: int) -> Any:
        raise NotImplementedError("Method '__getitem__' is not implemented.")

    @classmethod
    def prepare(cls, *args, **kwargs) -> None:
        raise NotImplementedError("Method 'prepare' is not implemented.")


__all__ = ["BaseDataset", "Split"]
