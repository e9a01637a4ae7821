"""Host-side logic on CPU: drop-in signatures, RNG replay, geometry, sampler sharding and the
metric exchange over a world_size-2 gloo group."""
import inspect
import os
import sys
from pathlib import Path

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from oracle import metrics_oracle as MO
from torchbooster_b200 import distributed as D
from torchbooster_b200 import metrics as M
from torchbooster_b200.transforms import BatchTransform, CropFlipParams, TransformSpec, resized_output_size

REFERENCE = Path("/root/reference")


# ---- drop-in surface --------------------------------------------------------------------------
@pytest.mark.skipif(not REFERENCE.exists(), reason="reference checkout only exists in the build container")
def test_signatures_match_reference():
    sys.path.insert(0, str(REFERENCE))
    try:
        import torchbooster.distributed as RD
        import torchbooster.metrics as RM
    finally:
        sys.path.remove(str(REFERENCE))

    def params(fn):
        return [(p.name, p.default, p.kind) for p in inspect.signature(fn).parameters.values()]

    for name in ("accuracy",):
        assert params(getattr(M, name)) == params(getattr(RM, name))
    assert params(M.Accuracy.forward) == params(RM.Accuracy.forward)
    assert params(M.RunningAverage.update) == params(RM.RunningAverage.update)
    for name in ("get_rank", "get_local_rank", "gather", "is_primary", "synchronize", "get_world_size",
                 "data_sampler", "find_free_port", "launch", "job"):
        assert params(getattr(D, name)) == params(getattr(RD, name)), name
    assert D.LOCAL_PROCESS_GROUP is None and RD.LOCAL_PROCESS_GROUP is None


def test_lmdb_and_dataset_signatures_match_reference_source():
    """py-lmdb is absent, so the reference modules cannot be imported: compare against their source."""
    if not REFERENCE.exists():
        pytest.skip("reference checkout only exists in the build container")
    import ast
    from torchbooster_b200 import dataset as DS
    from torchbooster_b200 import lmdb as LM

    def ref_methods(path, cls):
        tree = ast.parse((REFERENCE / path).read_text())
        node = next(n for n in tree.body if isinstance(n, ast.ClassDef) and n.name == cls)
        return {f.name: [a.arg for a in f.args.args] for f in node.body if isinstance(f, ast.FunctionDef)}

    for name, args in ref_methods("torchbooster/lmdb.py", "LMDBReader").items():
        assert [p for p in inspect.signature(getattr(LM.LMDBReader, name)).parameters][:len(args)] == args, name
    for name, args in ref_methods("torchbooster/dataset.py", "BaseDataset").items():
        got = [p for p in inspect.signature(getattr(DS.BaseDataset, name)).parameters]
        if name == "prepare":
            continue
        assert got[:len(args)] == args, name
    sig = inspect.signature(LM.LMDBReader.__init__).parameters
    assert sig["map_size"].default == 1024 ** 4 and sig["max_readers"].default == 126
    sig = inspect.signature(DS.BaseDataset.__init__).parameters
    assert sig["transform"].default is None and sig["map_size"].default == 1024 * 4 and sig["max_readers"].default == 126
    assert [s.value for s in DS.Split] == ["train", "validation", "test"]


def test_running_average_is_the_reference_recurrence(golden):
    g = golden("metrics_reference.npz")
    ra = M.RunningAverage()
    assert ra.current == 0 and ra.value == 0.
    for v, want in zip(g["ra_inputs"], g["ra_values"]):
        ra.update(float(v))
        assert ra.value == want
    assert ra.current == len(g["ra_inputs"])


def test_rank_helpers_without_process_group():
    assert D.get_rank() == 0 and D.get_local_rank() == 0 and D.get_world_size() == 1 and D.is_primary()
    D.synchronize()
    assert 0 < D.find_free_port() < 65536
    ran = []
    D.launch(lambda a, b: ran.append((a, b)), 1, args=(1, 2))
    assert ran == [(1, 2)]
    with pytest.raises(ValueError, match="auto"):
        D.launch(lambda: None, 1, n_machine=2, dist_url="auto")
    ds = list(range(10))
    assert type(D.data_sampler(ds, True, False)).__name__ == "RandomSampler"
    assert type(D.data_sampler(ds, False, False)).__name__ == "SequentialSampler"


@pytest.mark.skipif(torch.cuda.is_available(), reason="only meaningful without CUDA")
def test_job_requires_cuda():
    with pytest.raises(OSError, match="CUDA is not available"):
        D.job(0, lambda: None, 2, 2)


# ---- RNG replay and geometry --------------------------------------------------------------------
def test_crop_flip_replay_equals_torchvision_golden(golden):
    g = golden("params_torchvision.npz")
    for name in ("c2", "c1", "eq", "rect"):
        seed, B, h, w, th, tw, pad = g[f"{name}_cfg"].tolist()
        top, left, flip = CropFlipParams(seed).draw(B, h, w, (th, tw), (pad, pad), 0.5)
        want = g[f"{name}_params"]
        assert np.array_equal(top, want[:, 0]) and np.array_equal(left, want[:, 1]) and np.array_equal(flip, want[:, 2]), name


def test_crop_flip_replay_continues_across_batches(golden):
    g = golden("params_torchvision.npz")
    rng = CropFlipParams(42)
    a = rng.draw(40, 256, 256, (224, 224), (0, 0), 0.5)
    b = rng.draw(24, 256, 256, (224, 224), (0, 0), 0.5)
    got = np.stack([np.concatenate([a[i], b[i]]) for i in range(3)], 1)
    assert np.array_equal(got, g["c2_params"])
    with pytest.raises(ValueError, match="larger than input"):
        rng.draw(1, 10, 10, (12, 12))


def test_resized_output_size_is_torchvisions():
    assert resized_output_size(375, 500, 256) == (256, 341)
    assert resized_output_size(500, 375, 256) == (341, 256)
    assert resized_output_size(480, 640, 512) == (512, 682)
    assert resized_output_size(300, 200, (384, 256)) == (384, 256)
    assert resized_output_size(512, 512, [256]) == (256, 256)


def test_spec_validation():
    with pytest.raises(ValueError, match="Padding mode"):
        TransformSpec(padding=1, padding_mode="wrap")
    with pytest.raises(ValueError):
        TransformSpec(mean=[.5])
    with pytest.raises(ValueError, match="unsupported dtype"):
        TransformSpec(dtype=torch.float64)
    if not torch.cuda.is_available():
        with pytest.raises((RuntimeError, AssertionError)):
            BatchTransform(TransformSpec(), device="cpu")


def test_two_stage_pipelines_are_recognised():
    """Which Resize / CenterCrop pipelines need the uint8 hand-off to the augment kernel (two launches) and which stay one
    launch: constant padding after a bare resize reads as zeros outside the resized image in the resize kernel itself."""
    one = [TransformSpec(resize=64), TransformSpec(resize=64, center_crop=56), TransformSpec(resize=64, crop=48, padding=4),
           TransformSpec(crop=32, padding=4, padding_mode="reflect"), TransformSpec(crop=32, rand_augment=True),
           TransformSpec(random_resized_crop=32, rand_augment=True)]
    two = [TransformSpec(resize=64, crop=48, padding=4, padding_mode="reflect"),        # the padding mirrors the resized image
           TransformSpec(center_crop=40, crop=40, padding=4),                          # must not show what the centre crop cut
           TransformSpec(resize=64, center_crop=56, crop=56, padding=2, padding_mode="edge"),
           TransformSpec(resize=256, center_crop=224, flip_p=.5, rand_augment=True), TransformSpec(resize=64, rotation=10)]
    assert not any(sp.two_stage for sp in one)
    assert all(sp.two_stage for sp in two)
    with pytest.raises(ValueError, match="excludes"):
        TransformSpec(random_resized_crop=32, resize=64)


def test_cpu_inputs_are_rejected_not_emulated():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        M.accuracy(torch.randn(4, 3), torch.zeros(4, dtype=torch.int64))
    st = M.MetricState(3, k=2, device="cpu")          # buffers may live anywhere; compute may not
    with pytest.raises((RuntimeError, ValueError)):
        st.update(torch.randn(4, 3), torch.zeros(4, dtype=torch.int64))


# ---- world_size 2 over gloo ---------------------------------------------------------------------
def _worker(rank, world, port, tmp):
    D.BACKEND, D.REQUIRE_CUDA = "gloo", False
    out = {}

    def body():
        out["rank"], out["world"], out["local"] = D.get_rank(), D.get_world_size(), D.get_local_rank()
        # shard exactly like the reference: DistributedSampler(dataset, shuffle=True), seed 0, epoch 0
        ds = list(range(101))
        idx = list(D.data_sampler(ds, shuffle=False, distributed=True))
        out["indices"] = idx
        gen = torch.Generator().manual_seed(99)
        logits = torch.randn(101, 7, generator=gen)
        labels = torch.randint(0, 7, (101,), generator=gen)
        mine_l, mine_y = logits[idx], labels[idx]
        st = M.MetricState(7, k=3, device="cpu")
        rank_ = MO.label_rank(mine_l, mine_y)
        st.buffer[0] = len(idx)
        st.buffer[1] = int((rank_ < 1).sum())
        st.buffer[2] = int((rank_ < 3).sum())
        st.buffer[4:] = torch.from_numpy(MO.confusion(mine_l, mine_y, 7)).reshape(-1)
        st.loss[0], st.loss[1] = float(rank), 1.0
        st.all_reduce()
        D.synchronize()
        out["buffer"], out["loss"] = st.buffer.clone(), st.loss.clone()
        t = torch.tensor([float(D.get_rank())])
        lst = [torch.zeros(1) for _ in range(world)] if D.is_primary() else None
        D.gather(t, lst)
        out["gathered"] = None if lst is None else [float(x) for x in lst]

    D.job(rank, body, world, world, 0, f"tcp://127.0.0.1:{port}", ())
    torch.save(out, os.path.join(tmp, f"rank{rank}.pt"))
    dist.destroy_process_group()


def test_world_size_2_gloo(tmp_path):
    port = D.find_free_port()
    mp.spawn(_worker, nprocs=2, args=(2, port, str(tmp_path)), join=True)
    res = [torch.load(tmp_path / f"rank{r}.pt", weights_only=False) for r in range(2)]
    assert [r["rank"] for r in res] == [0, 1] and all(r["world"] == 2 for r in res)
    assert [r["local"] for r in res] == [0, 1]
    # shards are DistributedSampler's: randperm(101, seed 0), head-padded to 102, rank::2
    g = torch.Generator().manual_seed(0)
    perm = torch.randperm(101, generator=g).tolist()
    perm += perm[:1]
    assert res[0]["indices"] == perm[0::2] and res[1]["indices"] == perm[1::2]
    # reduced counters equal the single-process oracle over the concatenated shards
    gen = torch.Generator().manual_seed(99)
    logits = torch.randn(101, 7, generator=gen)
    labels = torch.randint(0, 7, (101,), generator=gen)
    allidx = res[0]["indices"] + res[1]["indices"]
    L, Y = logits[allidx], labels[allidx]
    for r in res:
        assert r["buffer"][0] == 102      # head padding duplicates one sample (inherent to the sampler)
        assert r["buffer"][1] == MO.topk_correct(L, Y, 1) and r["buffer"][2] == MO.topk_correct(L, Y, 3)
        assert np.array_equal(r["buffer"][4:].view(7, 7).numpy(), MO.confusion(L, Y, 7))
        assert r["loss"].tolist() == [1.0, 2.0]
    assert res[0]["gathered"] == [0.0, 1.0] and res[1]["gathered"] is None


def test_resized_crop_replay_equals_torchvision_golden(golden):
    from torchbooster_b200.transforms import ResizedCropParams
    g = golden("params_torchvision.npz")
    for name in ("rrc_c2", "rrc_wide", "rrc_tall"):
        cfg = g[f"{name}_cfg"]
        seed, B, h, w = (int(v) for v in cfg[:4])
        rng = ResizedCropParams(seed, scale=tuple(cfg[4:6]), ratio=tuple(cfg[6:8]))
        half = B // 3
        b1, f1 = rng.draw(half, h, w, 0.5)          # two calls: the unused words carry over
        b2, f2 = rng.draw(B - half, h, w, 0.5)
        want = g[f"{name}_params"]
        assert np.array_equal(np.concatenate([b1, b2]), want[:, :4]), name
        assert np.array_equal(np.concatenate([f1, f2]), want[:, 4]), name
    with pytest.raises(ValueError, match="excludes"):
        TransformSpec(random_resized_crop=24, crop=8)


def test_resized_crop_replay_over_images_of_different_sizes():
    """draw_ragged: every sample draws from its own size, in sequence, like T.RandomResizedCrop over a
    dataset of differently sized images; two calls continue one word stream."""
    import torchvision.transforms as T
    from PIL import Image
    from torchbooster_b200.transforms import ResizedCropParams
    rng = np.random.default_rng(8)
    hs, ws = rng.integers(8, 400, 300), rng.integers(8, 400, 300)
    state = torch.get_rng_state()
    try:
        torch.manual_seed(99)
        want, flips = [], []
        for h, w in zip(hs, ws):
            want.append(T.RandomResizedCrop.get_params(Image.new("L", (int(w), int(h))), [0.08, 1.0], [3 / 4, 4 / 3]))
            flips.append(int(torch.rand(1) < 0.5))
    finally:
        torch.set_rng_state(state)
    rp = ResizedCropParams(99)
    b1, f1 = rp.draw_ragged(hs[:120], ws[:120], 0.5)
    b2, f2 = rp.draw_ragged(hs[120:], ws[120:], 0.5)
    assert np.array_equal(np.concatenate([b1, b2]), np.array(want, np.int32))
    assert np.concatenate([f1, f2]).tolist() == flips


def test_bench_cpu_arm_staggers_the_workers():
    """bench.py's CPU arm: the first batch of every DataLoader worker is shorter, (w + 1) / workers of a batch, so that
    the lockstepped workers end up evenly out of phase; every later batch is a full one; CONFIGS name every BASELINE
    configuration with its algorithmic bytes."""
    import sys
    from pathlib import Path
    sys.path.insert(0, str(Path(__file__).resolve().parent.parent))
    import bench
    sizes = [len(b) for b in bench._StaggeredBatches(pool=100, B=64, workers=4, batches=7)]
    assert sizes == [16, 32, 48, 64, 64, 64, 64]
    assert all(0 <= i < 100 for b in bench._StaggeredBatches(100, 64, 4, 3) for i in b)
    assert {"c1", "c2", "c3", "c4", "c5", "c5_bf16"} <= set(bench.CONFIGS)
    assert bench.CONFIGS["c2"]["alg_bytes"] == 451584 and bench.CONFIGS["c4"]["alg_bytes"] == 1572864
    assert bench.CONFIGS["c5"]["alg_bytes"] == 4008 and bench.CONFIGS["c1"]["alg_bytes"] == 15360
