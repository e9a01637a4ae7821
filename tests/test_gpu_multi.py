"""Multi-GPU ring (needs >= 2 GPUs; skipped on a one-GPU box): ranks launched through the drop-in
`launch`, sharded by `data_sampler`, metrics summed by one NCCL all_reduce; the reduced state must
equal the single-process oracle over the concatenated shards."""
import os

import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO

pytestmark = pytest.mark.gpu


def _rank_main(tmp, n_classes, transport="int64"):
    from torchbooster_b200 import distributed as D
    from torchbooster_b200.metrics import MetricState
    rank, world = D.get_rank(), D.get_world_size()
    dev = torch.device("cuda", torch.cuda.current_device())
    assert D.get_local_rank() == rank and dev.index == rank
    n = 1001
    gen = torch.Generator().manual_seed(17)
    logits = torch.randn(n, n_classes, generator=gen)
    labels = torch.randint(0, n_classes, (n,), generator=gen)
    idx = list(D.data_sampler(list(range(n)), shuffle=True, distributed=True))
    st = MetricState(n_classes, k=5, device=dev)
    for s in range(0, len(idx), 128):
        part = idx[s:s + 128]
        st.update(logits[part].to(dev), labels[part].to(dev))
    st.update_loss(torch.full((len(idx),), float(rank + 1), device=dev))
    st.all_reduce(transport=transport)
    D.synchronize()
    res = st.compute()
    torch.save({"idx": idx, "buffer": st.buffer.cpu(), "loss": st.loss.cpu(), "n": res["n"]},
               os.path.join(tmp, f"rank{rank}.pt"))


@pytest.mark.skipif(torch.cuda.device_count() < 2, reason="needs at least 2 GPUs")
@pytest.mark.parametrize("transport", ["int64", "int32"])
def test_sharded_metrics_all_reduce_nccl(tmp_path, transport):
    from torchbooster_b200 import distributed as D
    world = min(torch.cuda.device_count(), 8)
    C = 37
    D.launch(_rank_main, world, dist_url="auto", args=(str(tmp_path), C, transport))
    res = [torch.load(tmp_path / f"rank{r}.pt", weights_only=False) for r in range(world)]
    gen = torch.Generator().manual_seed(17)
    logits = torch.randn(1001, C, generator=gen)
    labels = torch.randint(0, C, (1001,), generator=gen)
    allidx = sum((r["idx"] for r in res), [])
    per = -(-1001 // world)
    assert all(len(r["idx"]) == per for r in res) and len(set(allidx)) == 1001
    L, Y = logits[allidx], labels[allidx]
    for r in res:
        assert r["n"] == per * world
        assert r["buffer"][1] == MO.topk_correct(L, Y, 1) and r["buffer"][2] == MO.topk_correct(L, Y, 5)
        assert np.array_equal(r["buffer"][4:].view(C, C).numpy(), MO.confusion(L, Y, C))
        assert r["loss"][1] == per * world
        assert abs(float(r["loss"][0]) - per * sum(range(1, world + 1))) < 1e-9
