"""LMDBReader drop-in over this package's own reader/writer (py-lmdb is absent: the contract is
the behaviour cited from torchbooster/lmdb.py; parity against a py-lmdb-written file is UNPINNED)."""
import os

import numpy as np
import pytest

from torchbooster_b200.dataset import BaseDataset, Split
from torchbooster_b200.lmdb import LMDBReader, write_image_lmdb, write_lmdb


@pytest.fixture(scope="module")
def image_db(tmp_path_factory):
    d = tmp_path_factory.mktemp("db")
    rng = np.random.default_rng(0)
    imgs = rng.integers(0, 256, (200, 32, 32, 3), dtype=np.uint8)     # 3072-byte values: overflow pages
    write_image_lmdb(str(d / "cifar"), imgs, labels=np.arange(200) % 10)
    return str(d / "cifar"), imgs


def test_reader_contract(image_db):
    path, imgs = image_db
    r = LMDBReader(path)
    assert (r.path, r.map_size, r.max_readers, r.env, r.length) == (path, 1024 ** 4, 126, None, None)
    assert len(r) == 200 and r.env is None          # __len__ opens and closes (fork safety, lmdb.py:85-90)
    assert r[0] == imgs[0].tobytes() and r.env is not None     # lazy open on first get
    assert r.get(b"199") == imgs[199].tobytes()
    assert [len(v) for v in list(r)[:3]] == [3072] * 3
    with pytest.raises(KeyError, match="does not contain key"):
        r.get(b"200")
    r.close()
    assert r.env is None
    with LMDBReader(path) as ctx:
        assert ctx.get(b"label/13") == b"3"
    assert ctx.env is None


def test_missing_database_raises_ioerror(tmp_path):
    with pytest.raises(IOError, match="Could not open lmdb dataset"):
        LMDBReader(str(tmp_path / "absent")).open()
    (tmp_path / "junk").mkdir()
    (tmp_path / "junk" / "data.mdb").write_bytes(b"\0" * 8192)
    with pytest.raises(IOError):
        LMDBReader(str(tmp_path / "junk")).open()


def test_length_defaults_to_zero_without_length_key(tmp_path):
    write_lmdb(str(tmp_path / "nolen"), [(b"0", b"abc")])
    r = LMDBReader(str(tmp_path / "nolen"))
    r.open()
    assert r.length == 0 and r[0] == b"abc"


def test_small_values_deep_tree_and_key_order(tmp_path):
    items = [(f"{i}".encode(), f"value-{i}".encode() * (1 + i % 5)) for i in range(120_000)]
    items += [(b"length", b"120000"), (b"\x00\xff", b"lo"), (b"\xff\xfe", b"hi"), (b"k" * 511, b"longkey")]
    write_lmdb(str(tmp_path / "deep"), items)
    r = LMDBReader(str(tmp_path / "deep"))
    assert len(r) == 120_000
    for i in (0, 1, 9, 10, 99_999, 119_999, 54_321):
        assert r[i] == f"value-{i}".encode() * (1 + i % 5)
    assert r.get(b"\x00\xff") == b"lo" and r.get(b"\xff\xfe") == b"hi" and r.get(b"k" * 511) == b"longkey"
    for missing in (b"120000", b"", b"-1", b"k" * 510, b"\xff\xff"):
        with pytest.raises(KeyError):
            r.get(missing)


def test_empty_database(tmp_path):
    write_lmdb(str(tmp_path / "empty"), [])
    r = LMDBReader(str(tmp_path / "empty"))
    assert len(r) == 0
    with pytest.raises(KeyError):
        r.get(b"0")


def test_get_batch_gathers_records(image_db):
    path, imgs = image_db
    r = LMDBReader(path)
    idx = [5, 0, 199, 5, 42]
    out = r.get_batch(idx, 3072, threads=3)
    assert out.shape == (5, 3072) and np.array_equal(out.numpy().reshape(5, 32, 32, 3), imgs[idx])
    with pytest.raises(KeyError):
        r.get_batch([1, 200], 3072)
    with pytest.raises(ValueError, match="not 100 bytes"):
        r.get_batch([1], 100)
    assert r.get_batch([], 3072).shape == (0, 3072)


def test_reader_survives_fork(image_db):
    path, imgs = image_db
    r = LMDBReader(path)
    assert r[3] == imgs[3].tobytes()
    pid = os.fork()
    if pid == 0:
        ok = r[4] == imgs[4].tobytes()      # child reopens its own map
        os._exit(0 if ok else 1)
    assert os.waitpid(pid, 0)[1] == 0
    assert r[5] == imgs[5].tobytes()


def test_base_dataset_contract(image_db):
    path, imgs = image_db
    ds = BaseDataset(path)
    assert (ds.path, ds.transform, ds.map_size, ds.max_readers) == (path, None, 1024 * 4, 126)
    assert isinstance(ds.lmdb_reader, LMDBReader) and len(ds) == 200
    with pytest.raises(NotImplementedError, match="__getitem__"):
        ds[0]
    with pytest.raises(NotImplementedError, match="prepare"):
        BaseDataset.prepare()
    assert Split("validation") is Split.VALID

    class Cifar(BaseDataset):
        def __getitem__(self, idx):
            img = np.frombuffer(self.lmdb_reader[idx], np.uint8).reshape(32, 32, 3)
            return img, int(self.lmdb_reader.get(f"label/{idx}".encode()))

    x, y = Cifar(path)[17]
    assert np.array_equal(x, imgs[17]) and y == 7


def test_batch_loader_host_side(image_db):
    """Order, ragged last batch, drop_last, labels, prefetch and early exit -- without a GPU."""
    from torchbooster_b200.loader import BatchLoader
    path, imgs = image_db
    reader = LMDBReader(path)
    labels = reader.read_labels()
    assert labels.tolist() == (np.arange(200) % 10).tolist()
    order = np.random.default_rng(3).permutation(200)[:77]
    loader = BatchLoader(reader, (32, 32, 3), batch_size=16, sampler=order.tolist(), labels=labels, prefetch=3, threads=2)
    assert len(loader) == 5
    seen = list(loader)
    assert [x.shape[0] for x, _ in seen] == [16, 16, 16, 16, 13]
    got = np.concatenate([x.numpy() for x, _ in seen])
    assert np.array_equal(got, imgs[order])
    assert torch_equal_list([y for _, y in seen], labels[order])
    dl = BatchLoader(reader, (32, 32, 3), batch_size=16, sampler=order.tolist(), drop_last=True)
    assert len(dl) == 4 and sum(x.shape[0] for x, y in dl) == 64
    full = BatchLoader(reader, (32, 32, 3), batch_size=64)
    assert len(full) == 4 and sum(x.shape[0] for x, _ in full) == 200
    it = iter(BatchLoader(reader, (32, 32, 3), batch_size=8, prefetch=1))
    next(it); next(it)
    it.close()                                    # early exit must not hang on the producer
    bad = BatchLoader(reader, (32, 32, 3), batch_size=8, sampler=[1, 2, 999])
    with pytest.raises(KeyError):
        list(bad)


def torch_equal_list(tensors, want):
    import torch
    return np.array_equal(torch.cat(tensors).numpy(), np.asarray(want))


def test_ragged_image_records_round_trip(tmp_path):
    """Images of different sizes: 16-byte TBI1 header + rows padded to 4 bytes; gathered back to back with
    16-byte aligned offsets; missing keys and a reused (too small, then large enough) buffer."""
    import torch
    from torchbooster_b200.lmdb import RAGGED_MAGIC, write_ragged_image_lmdb
    rng = np.random.default_rng(2)
    sizes = [(10, 7), (33, 21), (5, 40), (1, 1), (64, 50), (300, 200)]          # the last ones live on overflow pages
    imgs = [rng.integers(0, 256, (h, w, 3), dtype=np.uint8) for h, w in sizes]
    write_ragged_image_lmdb(str(tmp_path / "db"), imgs, labels=range(len(imgs)))
    reader = LMDBReader(str(tmp_path / "db"))
    assert len(reader) == len(imgs) and reader.read_labels().tolist() == list(range(len(imgs)))
    order = [5, 0, 3, 1, 4, 2]
    small = torch.empty(64, dtype=torch.uint8)
    buf, off, sz = reader.get_ragged_batch(order, out=small, threads=3)
    assert buf.numel() >= int(off[-1] + sz[-1]) and (off % 16 == 0).all()
    again, off2, _ = reader.get_ragged_batch(order, out=buf, threads=1)
    assert again.data_ptr() == buf.data_ptr() and np.array_equal(off, off2)
    hv = buf.numpy()
    for o, i in zip(off, order):
        magic, h, w, c = (int(v) for v in hv[o:o + 16].view("<u4"))
        assert magic == RAGGED_MAGIC and (h, w, c) == imgs[i].shape
        pitch = (w * c + 3) // 4 * 4
        assert np.array_equal(hv[o + 16:o + 16 + h * pitch].reshape(h, pitch)[:, :w * c].reshape(h, w, c), imgs[i])
    with pytest.raises(KeyError):
        reader.get_ragged_batch([0, 99])
    assert reader.get_ragged_batch([])[1].shape == (0,)


def test_reopen_and_len_do_not_leak_or_close_an_open_reader(tmp_path):
    """open() twice releases the first handle; len() on an OPEN reader whose database has no length key (length
    0) neither reopens nor closes it (only a closed reader does the reference's open-and-close, lmdb.py:85-90)."""
    write_lmdb(str(tmp_path / "nolen"), [(b"0", b"abc")])
    r = LMDBReader(str(tmp_path / "nolen"))
    r.open()
    first = r.env.value
    r.open()
    assert r.env is not None and r.get(b"0") == b"abc"
    held = r.env.value
    assert len(r) == 0 and r.env is not None and r.env.value == held
    r.close()
    assert len(r) == 0 and r.env is None
    del first


def test_unsupported_main_db_flags_are_refused(tmp_path):
    """A main database with reverse / integer keys or duplicates is not misread as 'key not found': it does not open."""
    write_lmdb(str(tmp_path / "db"), [(b"0", b"abc"), (b"length", b"1")])
    f = tmp_path / "db" / "data.mdb"
    raw = bytearray(f.read_bytes())
    psize = int.from_bytes(raw[16 + 24:16 + 28], "little")
    for page in (0, 1):
        off = page * psize + 16 + 24 + 48 + 4           # meta: dbs[1].md_flags
        raw[off:off + 2] = (0x08).to_bytes(2, "little")  # MDB_INTEGERKEY
    f.write_bytes(bytes(raw))
    with pytest.raises(IOError, match="not supported"):
        LMDBReader(str(tmp_path / "db")).open()


def test_map_is_read_only_and_unreserved(tmp_path):
    """The reader's map is PROT_READ (what readonly=True gives the reference): it shows as 'r--p' in /proc/self/maps."""
    write_lmdb(str(tmp_path / "db"), [(b"0", b"abc"), (b"length", b"1")])
    r = LMDBReader(str(tmp_path / "db"))
    r.open()
    lines = [ln for ln in open("/proc/self/maps") if str(tmp_path / "db" / "data.mdb") in ln]
    assert lines and all(ln.split()[1] == "r--p" for ln in lines)
    r.close()
