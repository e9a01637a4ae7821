"""LMDB byte source -- drop-in for ``torchbooster/lmdb.py`` that needs no py-lmdb.

``LMDBReader`` keeps the reference's surface (``torchbooster/lmdb.py:13-106``): constructor
``(path, map_size=1024**4, max_readers=126)``, attributes ``path map_size max_readers env
length``, lazy per-process ``open()``, ``get(key) -> bytes`` raising ``KeyError`` for a missing
key, ``length`` read from the ASCII ``b"length"`` record (0 when absent), ``__len__`` that opens and
closes to stay fork-safe, ``__getitem__(idx) = get(str(idx).encode())``.

The environment is this library's own read-only reader (``tb_lmdb_*`` in the C ABI: mmap of
``data.mdb``, B-tree descent, no locks -- the reference opens with ``lock=False`` too).
``map_size`` and ``max_readers`` are accepted and ignored: a read-only private mapping needs neither.

Additions: ``get_batch`` (many records -> one pinned uint8 tensor, the input of
``BatchTransform``) and ``write_lmdb`` (emits the same on-disk subset; used by ``prepare()``
implementations, tests and the benchmark since no LMDB writer exists on the box).
"""
from __future__ import annotations

import ctypes as C
import os
import struct
from typing import Iterable, Iterator, Optional, Tuple

from . import _native as N


class LMDBReader:
    """LMDB Reader.

    The database must hold a ``length`` key whose value is the dataset size in ASCII.

    Parameters
    ----------
    path: str
        path to the lmdb dataset (directory holding ``data.mdb``, or the data file)
    map_size, max_readers: int (default: 1024 ** 4, 126)
        kept for signature compatibility; unused by the read-only reader

    Attributes
    ----------
    env: opaque handle of the open environment, ``None`` while closed
    length: int dataset size (``None`` until first opened)
    """

    def __init__(self, path: str, map_size: int = 1024 ** 4, max_readers: int = 126) -> None:
        self.path = path
        self.map_size = map_size
        self.max_readers = max_readers

        self.env = None
        self.length = None

    def open(self) -> None:
        """Open the environment (per process; never carried across fork).  A handle this process already holds is
        released first, so opening twice does not leak the previous map."""
        self.close()
        handle = C.c_void_p()
        rc = N.lib().tb_lmdb_open(os.fsencode(str(self.path)), C.byref(handle))
        if rc != 0 or not handle:
            raise IOError(f"Could not open lmdb dataset {self.path}: {N.last_error()}")
        self.env = handle
        self._pid = os.getpid()
        try:
            self.length = int(self.get(b"length").decode("utf-8"))
        except KeyError:
            self.length = 0

    def close(self) -> None:
        if self.env is not None:
            if getattr(self, "_pid", None) == os.getpid():
                N.lib().tb_lmdb_close(self.env)
            self.env = None

    def get(self, key: bytes) -> bytes:
        """Value stored under ``key`` (a copy out of the map); ``KeyError`` if absent."""
        if self.env is None or getattr(self, "_pid", None) != os.getpid():
            self.env = None
            self.open()
        data, size = C.c_void_p(), C.c_int64()
        rc = N.lib().tb_lmdb_get(self.env, key, len(key), C.byref(data), C.byref(size))
        if rc == N.TB_ENOENT:
            raise KeyError(f"lmdb dataset does not contain key {key}")
        N.check(rc)
        return C.string_at(data, size.value)

    def get_batch(self, indices, record_bytes: int, out=None, threads: int = 8):
        """Records ``str(i)`` for ``i`` in ``indices`` -> one ``uint8 [len(indices), record_bytes]``
        pinned host tensor (the uint8 batch ``BatchTransform`` consumes), gathered by
        ``threads`` host threads straight out of the map."""
        import numpy as np
        import torch
        if self.env is None or getattr(self, "_pid", None) != os.getpid():
            self.env = None
            self.open()
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64))
        n = int(idx.shape[0])
        if out is None:
            out = torch.empty((n, record_bytes), dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        elif out.dtype != torch.uint8 or out.numel() != n * record_bytes or not out.is_contiguous() or out.is_cuda:
            raise ValueError("out must be a contiguous host uint8 tensor of n * record_bytes elements")
        rc = N.lib().tb_lmdb_gather(self.env, idx.ctypes.data, n, record_bytes, out.data_ptr(), threads)
        if rc == N.TB_ENOENT:
            raise KeyError(N.last_error())
        N.check(rc)
        return out

    def get_ragged_batch(self, indices, out=None, threads: int = 8):
        """Records ``str(i)`` of DIFFERENT sizes -> ``(buffer, offsets, sizes)``: one pinned ``uint8`` host tensor
        holding them back to back (every record starts on a 16-byte boundary), the byte offset and length of
        each.  ``out``: a pinned tensor to reuse when it is large enough."""
        import numpy as np
        import torch
        if self.env is None or getattr(self, "_pid", None) != os.getpid():
            self.env = None
            self.open()
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64))
        n = int(idx.shape[0])
        sizes = np.zeros(n, np.int64)
        L = N.lib()

        def call(offsets, dst, capacity):
            rc = L.tb_lmdb_gather_ragged(self.env, idx.ctypes.data, n, sizes.ctypes.data,
                                         None if offsets is None else offsets.ctypes.data, dst, capacity, threads)
            if rc == N.TB_ENOENT:
                raise KeyError(N.last_error())
            N.check(rc)

        call(None, None, 0)
        padded = (sizes + 15) // 16 * 16
        offsets = np.concatenate([[0], np.cumsum(padded)[:-1]]).astype(np.int64) if n else np.zeros(0, np.int64)
        total = int(padded.sum())
        if out is None or out.numel() < total:
            out = torch.empty(max(total, 16), dtype=torch.uint8, pin_memory=torch.cuda.is_available())
        elif out.dtype != torch.uint8 or not out.is_contiguous() or out.is_cuda:
            raise ValueError("out must be a contiguous host uint8 tensor")
        call(offsets, out.data_ptr(), out.numel())
        return out, offsets, sizes

    def cuda_register(self) -> int:
        """Page-lock the map and map it into the GPU's address space; returns the device address
        of its first byte.  Afterwards ``BatchTransform.from_records`` / ``BatchLoader(in_place=True)``
        let the kernel read records where they lie (no gather into a staging buffer).  The whole
        database becomes resident, so this is for databases that fit in host memory.  Raises
        ``NotImplementedError`` when the platform cannot register the mapping."""
        if self.env is None or getattr(self, "_pid", None) != os.getpid():
            self.env = None
            self.open()
        base = C.c_void_p()
        N.check(N.lib().tb_lmdb_cuda_register(self.env, C.byref(base)))
        return int(base.value)

    def record_offsets(self, indices, record_bytes: int):
        """Byte offsets (from the start of the map) of the records ``str(i)``, as numpy int64."""
        import numpy as np
        if self.env is None or getattr(self, "_pid", None) != os.getpid():
            self.env = None
            self.open()
        idx = np.ascontiguousarray(np.asarray(indices, dtype=np.int64))
        off = np.empty(idx.shape[0], np.int64)
        rc = N.lib().tb_lmdb_record_offsets(self.env, idx.ctypes.data, idx.shape[0], record_bytes, off.ctypes.data)
        if rc == N.TB_ENOENT:
            raise KeyError(N.last_error())
        N.check(rc)
        return off

    def read_labels(self):
        """Label array written by ``write_image_lmdb`` (record ``labels`` = raw little-endian int64),
        or ``None`` when the database has none."""
        import numpy as np
        try:
            return np.frombuffer(self.get(b"labels"), dtype="<i8").copy()
        except KeyError:
            return None

    def __len__(self) -> int:
        # torchbooster/lmdb.py:85-90: a reader that is still closed opens and closes again so that no
        # environment is carried across a later fork; a reader already open in this process stays open.
        if not self.length:
            was_open = self.env is not None and getattr(self, "_pid", None) == os.getpid()
            if not was_open:
                self.open()
                self.close()
        return self.length

    def __iter__(self) -> Iterator[bytes]:
        for i in range(len(self)):
            yield self[i]

    def __getitem__(self, idx: int) -> bytes:
        return self.get(str(idx).encode("utf-8"))

    def __del__(self) -> None:
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self) -> "LMDBReader":
        return self

    def __exit__(self, *args, **kwargs) -> None:
        self.close()


# ---------------------------------------------------------------------------------------------
# writer (subset of the format the reader understands)
# ---------------------------------------------------------------------------------------------
_PSIZE = 4096
_HDR = 16
_NODEMAX = (((_PSIZE - _HDR) // 2) & ~1) - 2     # liblmdb: largest node kept inside a leaf page
_MAXKEY = 511


def _page(pgno: int, flags: int, lower: int, upper: int, body: bytes) -> bytes:
    head = struct.pack("<QHHHH", pgno, 0, flags, lower, upper)
    return (head + body).ljust(_PSIZE, b"\0")


def _build_pages(nodes, flags, next_pgno, pages):
    """Pack ``nodes`` = [(key, payload_bytes)] into pages of ``flags`` kind.  Returns
    [(first_key, pgno)] of the pages written."""
    out, cur, cur_keys, used = [], [], [], 0

    def flush():
        nonlocal cur, cur_keys, used
        if not cur:
            return
        pgno = next_pgno[0]
        next_pgno[0] += 1
        ptrs, blob, upper = [], bytearray(_PSIZE), _PSIZE
        for nd in cur:
            size = (len(nd) + 1) & ~1
            upper -= size
            blob[upper:upper + len(nd)] = nd
            ptrs.append(upper)
        lower = _HDR + 2 * len(cur)
        body = struct.pack(f"<{len(ptrs)}H", *ptrs) + bytes(blob[_HDR + 2 * len(ptrs):])
        pages[pgno] = _page(pgno, flags, lower, upper, body)
        out.append((cur_keys[0], pgno))
        cur, cur_keys, used = [], [], 0

    for key, node in nodes:
        need = ((len(node) + 1) & ~1) + 2
        if used + need > _PSIZE - _HDR:
            flush()
        cur.append(node)
        cur_keys.append(key)
        used += need
    flush()
    return out


def write_lmdb(path: str, items: Iterable[Tuple[bytes, bytes]], subdir: bool = True) -> str:
    """Write a fresh single-transaction LMDB environment holding ``items`` (key, value) pairs.

    Emits liblmdb's data format version 1 with 4 KiB pages: two meta pages, leaf pages, overflow
    runs for values that do not fit a leaf node, and as many branch levels as needed.  Keys are
    stored in memcmp order (LMDB's default comparator).  Returns the data file path.
    """
    pairs = sorted(((bytes(k), bytes(v)) for k, v in items), key=lambda kv: kv[0])
    for i, (k, _) in enumerate(pairs):
        if not 0 < len(k) <= _MAXKEY:
            raise ValueError("keys must be 1..511 bytes")
        if i and pairs[i - 1][0] == k:
            raise ValueError(f"duplicate key {k!r}")
    pages = {}
    next_pgno = [2]
    leaf_nodes, n_overflow = [], 0
    for k, v in pairs:
        if 8 + len(k) + len(v) > _NODEMAX:
            npages = (_HDR + len(v) + _PSIZE - 1) // _PSIZE
            pgno = next_pgno[0]
            next_pgno[0] += npages
            run = struct.pack("<QHHI", pgno, 0, 0x04, npages) + v
            run = run.ljust(npages * _PSIZE, b"\0")
            for j in range(npages):
                pages[pgno + j] = run[j * _PSIZE:(j + 1) * _PSIZE]
            n_overflow += npages
            node = struct.pack("<HHHH", len(v) & 0xFFFF, len(v) >> 16, 0x01, len(k)) + k + struct.pack("<Q", pgno)
        else:
            node = struct.pack("<HHHH", len(v) & 0xFFFF, len(v) >> 16, 0, len(k)) + k + v
        leaf_nodes.append((k, node))
    level = _build_pages(leaf_nodes, 0x02, next_pgno, pages)
    n_leaf, n_branch, depth = len(level), 0, 1 if level else 0
    while len(level) > 1:
        nodes = []
        # liblmdb writes branch pages whose node 0 has an empty key; page boundaries are decided
        # with the full key, then the first key of every page is dropped
        built, cur_first = [], True
        for key, pgno in level:
            nodes.append((key, (key, pgno)))
        packed, cur, used = [], [], 0
        for key, (k2, pgno) in nodes:
            need = ((8 + len(k2) + 1) & ~1) + 2
            if cur and used + need > _PSIZE - _HDR:
                packed.append(cur)
                cur, used = [], 0
            cur.append((k2, pgno))
            used += need
        if cur:
            packed.append(cur)
        new_level = []
        for group in packed:
            gnodes = []
            for j, (k2, pgno) in enumerate(group):
                kk = b"" if j == 0 else k2
                gnodes.append((k2, struct.pack("<HHHH", pgno & 0xFFFF, (pgno >> 16) & 0xFFFF, (pgno >> 32) & 0xFFFF,
                                               len(kk)) + kk))
            new_level += _build_pages(gnodes, 0x01, next_pgno, pages)
            new_level[-1] = (group[0][0], new_level[-1][1])
        n_branch += len(new_level)
        level = new_level
        depth += 1
    root = level[0][1] if level else (1 << 64) - 1
    last_pg = next_pgno[0] - 1

    def meta(pgno, txnid):
        free_db = struct.pack("<IHHQQQQQ", _PSIZE, 0, 0, 0, 0, 0, 0, (1 << 64) - 1)
        main_db = struct.pack("<IHHQQQQQ", 0, 0, depth, n_branch, n_leaf, n_overflow, len(pairs), root)
        body = struct.pack("<IIQQ", 0xBEEFC0DE, 1, 0, max(1 << 20, (last_pg + 1) * _PSIZE)) + free_db + main_db + \
            struct.pack("<QQ", last_pg if last_pg >= 1 else 1, txnid)
        return (struct.pack("<QHHHH", pgno, 0, 0x08, 0, 0) + body).ljust(_PSIZE, b"\0")

    if subdir:
        os.makedirs(path, exist_ok=True)
        file = os.path.join(path, "data.mdb")
    else:
        file = path
    with open(file, "wb") as f:
        f.write(meta(0, 0))
        f.write(meta(1, 1))
        for pgno in range(2, last_pg + 1):
            f.write(pages[pgno])
    return file


def write_image_lmdb(path: str, images, labels: Optional[Iterable[int]] = None) -> str:
    """Value encoding used by this package (the reference leaves it open): record
    ``str(i)`` = raw HWC uint8 bytes of image ``i``, ``label/<i>`` = ASCII label, ``labels`` = all
    labels as raw little-endian int64, ``length`` = count, ``shape`` = ASCII ``H,W,C``."""
    import numpy as np
    images = np.asarray(images)
    if images.dtype != np.uint8 or images.ndim != 4:
        raise ValueError("images must be uint8 [N,H,W,C]")
    items = [(str(i).encode(), images[i].tobytes()) for i in range(images.shape[0])]
    items.append((b"length", str(images.shape[0]).encode()))
    items.append((b"shape", ",".join(str(s) for s in images.shape[1:]).encode()))
    if labels is not None:
        labels = [int(l) for l in labels]
        items += [(f"label/{i}".encode(), str(l).encode()) for i, l in enumerate(labels)]
        items.append((b"labels", np.asarray(labels, dtype="<i8").tobytes()))
    return write_lmdb(path, items)


RAGGED_MAGIC = 0x31494254          # "TBI1": header of an image record that carries its own shape


def write_ragged_image_lmdb(path: str, images, labels: Optional[Iterable[int]] = None) -> str:
    """Images of DIFFERENT sizes: record ``str(i)`` = 16-byte header ``{magic "TBI1", H, W, C}`` (little-endian
    uint32) + the rows of the uint8 HWC image, each padded to a multiple of 4 bytes -- the layout
    ``BatchTransform.ragged_records`` hands to the kernel as it is.  ``labels`` / ``length`` as in
    ``write_image_lmdb``."""
    import numpy as np
    items = []
    for i, im in enumerate(images):
        a = np.ascontiguousarray(np.asarray(im))
        if a.dtype != np.uint8 or a.ndim != 3:
            raise ValueError("images must be uint8 [H,W,C] arrays")
        h, w, c = a.shape
        pitch = (w * c + 3) // 4 * 4
        rows = np.zeros((h, pitch), np.uint8)
        rows[:, :w * c] = a.reshape(h, w * c)
        items.append((str(i).encode(), np.array([RAGGED_MAGIC, h, w, c], "<u4").tobytes() + rows.tobytes()))
    n = len(items)
    items.append((b"length", str(n).encode()))
    items.append((b"shape", b"ragged"))
    if labels is not None:
        labels = [int(l) for l in labels]
        items += [(f"label/{i}".encode(), str(l).encode()) for i, l in enumerate(labels)]
        items.append((b"labels", np.asarray(labels, dtype="<i8").tobytes()))
    return write_lmdb(path, items)
