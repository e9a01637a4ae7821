"""Parity of the CUDA metrics path: counts and confusion bins are bit-exact against the
reference-pinned oracle; the accuracy quotient is the reference's float32 value exactly."""
import numpy as np
import pytest
import torch

from oracle import metrics_oracle as MO
from torchbooster_b200 import metrics as M

pytestmark = pytest.mark.gpu


def bf16_from_bits(bits):
    return torch.from_numpy(np.ascontiguousarray(bits)).view(torch.bfloat16)


def test_reference_golden(golden):
    g = golden("metrics_reference.npz")
    kat = torch.from_numpy(g["kat_logits"]).cuda()
    for lab in range(4):
        labels = torch.full((kat.shape[0],), lab, dtype=torch.int64, device="cuda")
        acc = M.accuracy(kat, labels)
        assert acc.dtype == torch.float32 and acc.dim() == 0 and acc.is_cuda
        assert acc.cpu().numpy() == g[f"kat_accuracy_label{lab}"]
    cm = M.confusion_matrix(kat, torch.zeros(kat.shape[0], dtype=torch.int64, device="cuda"))
    assert cm[0].cpu().tolist() == np.bincount(g["kat_argmax"], minlength=4).tolist()
    a = M.accuracy(torch.from_numpy(g["a_logits"]).cuda(), torch.from_numpy(g["a_labels"]).cuda())
    assert a.cpu().numpy() == g["a_accuracy"]
    b = M.accuracy(bf16_from_bits(g["b_logits_bf16_bits"]).cuda(), torch.from_numpy(g["b_labels"]).cuda())
    assert b.cpu().numpy() == g["b_accuracy"]
    c = M.Accuracy()(torch.from_numpy(g["c_logits"]).cuda(), torch.from_numpy(g["c_labels"]).cuda())
    assert c.cpu().numpy() == g["c_accuracy"] == 5.0
    d = M.accuracy(torch.from_numpy(g["d_logits"]).cuda(), torch.from_numpy(g["d_labels"]).cuda(), dim=0)
    assert d.cpu().numpy() == g["d_accuracy_dim0"]


def check_state(logits, labels, k, C):
    st = M.MetricState(C, k=k, device="cuda")
    st.update(logits.cuda(), labels.cuda())
    res = st.compute()
    rank = MO.label_rank(logits.float(), labels)
    valid = ((labels >= 0) & (labels < C)).numpy()
    assert res["n"] == logits.shape[0]
    assert res["correct_1"] == int(((rank < 1) & valid).sum())
    assert res[f"correct_{k}"] == int(((rank < k) & valid).sum())
    assert res["invalid_labels"] == int((~valid).sum())
    assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(logits.float(), labels, C))
    return res


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("N,C", [(1, 1000), (1000, 1000), (513, 10), (64, 7), (33, 1), (100, 1003), (257, 4099)])
def test_counts_and_confusion(dtype, N, C):
    gen = torch.Generator().manual_seed(N * 31 + C)
    logits = torch.randn(N, C, generator=gen).to(dtype)      # 16-bit values tie often: stresses the rank rule
    labels = torch.randint(0, C, (N,), generator=gen)
    check_state(logits, labels, 5, C)


def test_ties_nan_inf_signed_zero_and_invalid_labels():
    nan, inf = float("nan"), float("inf")
    neg_nan = nan   # replaced below by a negative NaN with a payload, bit for bit
    rows = torch.tensor([[2., 2, 2, 2, 2], [nan, 5, 1, nan, 0], [1, nan, 0, 0, 0], [-0.0, 0.0, -1, -2, -3],
                         [0.0, -0.0, -1, -2, -3], [-inf] * 5, [inf, nan, inf, 1, 0], [1, 2, 3, 4, 5],
                         [neg_nan, 9, 9, 9, 9], [3, neg_nan, nan, 3, 3]])
    bits = rows.view(torch.int32)
    bits[8, 0] = int(np.array([0xFFC00001], np.uint32).view(np.int32)[0])
    bits[9, 1] = int(np.array([0xFF800123], np.uint32).view(np.int32)[0])
    bits[9, 2] = 0x7F800001
    rows = torch.cat([rows, rows, rows])
    labels = torch.cat([torch.arange(10) % 5, (torch.arange(10) + 2) % 5, torch.tensor([5, -1, 0, 1, 2, 3, 4, 99, 0, 1])])
    for dtype in (torch.float32, torch.bfloat16, torch.float16):
        check_state(rows.to(dtype), labels, 2, 5)


def test_unaligned_rows_and_strided_view():
    gen = torch.Generator().manual_seed(4)
    base = torch.randn(50, 40, generator=gen)
    for dtype in (torch.float32, torch.bfloat16):
        big = base.to(dtype).cuda()
        view = big[:, 3:36]                      # row stride 40, 33 classes, rows start off 16-byte boundaries
        labels = torch.randint(0, 33, (50,), generator=gen)
        st = M.MetricState(33, k=3, device="cuda")
        st.update(view, labels.cuda())
        res = st.compute()
        rank = MO.label_rank(view.float().cpu(), labels)
        assert res["correct_1"] == int((rank < 1).sum()) and res["correct_3"] == int((rank < 3).sum())
        assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(view.float().cpu(), labels, 33))


def test_accumulation_over_batches_and_loss():
    gen = torch.Generator().manual_seed(8)
    st = M.MetricState(10, k=3, device="cuda")
    dra = M.DeviceRunningAverage("cuda")
    all_l, all_y, losses = [], [], []
    for _ in range(5):
        l = torch.randn(97, 10, generator=gen); y = torch.randint(0, 10, (97,), generator=gen)
        loss = torch.rand((), generator=gen)
        st.update(l.cuda(), y.cuda()); st.update_loss(loss.cuda()); dra.update(loss.cuda())
        all_l.append(l); all_y.append(y); losses.append(float(loss))
    res = st.compute()
    L, Y = torch.cat(all_l), torch.cat(all_y)
    assert res["n"] == 485 and res["correct_1"] == MO.topk_correct(L, Y, 1) and res["correct_3"] == MO.topk_correct(L, Y, 3)
    assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(L, Y, 10))
    assert abs(res["loss"] - float(np.mean(np.array(losses, np.float64)))) < 1e-12
    assert dra.current == 5 and abs(dra.value - res["loss"]) < 1e-15
    per_sample = torch.rand(1000, generator=gen)
    st.reset(); st.update_loss(per_sample.cuda())
    assert abs(st.compute()["loss"] - float(per_sample.double().mean())) < 1e-12


def test_topk_matches_torch_topk_without_ties():
    gen = torch.Generator().manual_seed(1)
    logits = torch.randn(4096, 1000, generator=gen)
    labels = torch.randint(0, 1000, (4096,), generator=gen)
    want = (logits.topk(5, -1).indices == labels[:, None]).any(-1).float().mean()
    got = M.top_k_accuracy(logits.cuda(), labels.cuda(), k=5)
    assert got.cpu() == want


def test_full_size_c5_properties():
    """BASELINE config 5 scale (1M x 1000 would be 4 GB; per-rank share 125k x 1000 here): the
    confusion matrix row sums equal the label histogram, its trace equals correct_1, and
    permuting the samples leaves every accumulator unchanged."""
    N, C = 125_000, 1000
    gen = torch.Generator(device="cuda").manual_seed(1234)
    logits = torch.randn(N, C, device="cuda", generator=gen)
    labels = torch.randint(0, C, (N,), device="cuda", generator=gen)
    labels[::3] = logits[::3].argmax(-1)
    st = M.MetricState(C, k=5, device="cuda")
    st.update(logits, labels)
    res = st.compute()
    cm = res["confusion"]
    assert res["n"] == N and int(cm.sum()) == N
    assert torch.equal(cm.sum(1), torch.bincount(labels, minlength=C))
    assert int(cm.diagonal().sum()) == res["correct_1"] == int((logits.argmax(-1) == labels).sum())
    assert res["correct_5"] >= res["correct_1"]
    perm = torch.randperm(N, device="cuda", generator=gen)
    st2 = M.MetricState(C, k=5, device="cuda")
    st2.update(logits[perm], labels[perm])
    assert torch.equal(st2.buffer, st.buffer)
    assert float(M.accuracy(logits, labels)) == float(torch.tensor(res["correct_1"]) / N)


def test_cpu_tensors_are_rejected():
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        M.accuracy(torch.randn(4, 3), torch.zeros(4, dtype=torch.int64))


@pytest.mark.parametrize("seed", range(4))
def test_fuzz_rows_against_rank_rule(seed):
    """Random widths, strides, dtypes, duplicated values (ties), NaN / inf sprinkles and invalid
    labels against the rank-rule oracle."""
    gen = torch.Generator().manual_seed(500 + seed)
    for _ in range(8):
        C = int(torch.randint(1, 300, (1,), generator=gen))
        n = int(torch.randint(1, 80, (1,), generator=gen))
        k = int(torch.randint(1, 7, (1,), generator=gen))
        dtype = [torch.float32, torch.bfloat16, torch.float16][int(torch.randint(0, 3, (1,), generator=gen))]
        pad = int(torch.randint(0, 9, (1,), generator=gen))
        base = torch.randint(-4, 5, (n, C + pad), generator=gen).float() / 2            # heavy ties
        mask = torch.rand(n, C + pad, generator=gen)
        base[mask < 0.02] = float("nan")
        base[(mask >= 0.02) & (mask < 0.03)] = float("inf")
        base[(mask >= 0.03) & (mask < 0.04)] = float("-inf")
        base[(mask >= 0.04) & (mask < 0.06)] = -0.0
        logits = base.to(dtype).cuda()[:, pad // 2: pad // 2 + C]                        # strided, often misaligned rows
        labels = torch.randint(-1, C + 1, (n,), generator=gen)
        st = M.MetricState(C, k=k, device="cuda")
        st.update(logits, labels.cuda())
        res = st.compute()
        ref = logits.float().cpu()
        rank = MO.label_rank(ref, labels)
        valid = ((labels >= 0) & (labels < C)).numpy()
        assert res["correct_1"] == int(((rank < 1) & valid).sum()), (seed, C, n, dtype)
        assert res[f"correct_{k}"] == int(((rank < k) & valid).sum()), (seed, C, n, k, dtype)
        assert res["invalid_labels"] == int((~valid).sum())
        assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(ref, labels, C)), (seed, C, n, dtype)


def test_int32_wire_format_sums_exactly():
    """The int32 transport of the exchange: three ranks' states packed, their wire buffers added as int32 (what the
    all_reduce does), unpacked -- equal to the int64 sum, header values far beyond 32 bits included."""
    import torch
    from torchbooster_b200 import _native as N
    C = 23
    gen = torch.Generator().manual_seed(5)
    states = []
    for r in range(3):
        st = M.MetricState(C, k=5, device="cuda")
        st.buffer[4:] = torch.randint(0, 2 ** 31 // 3, (C * C,), generator=gen).cuda()
        st.buffer[:4] = torch.tensor([2 ** 61 // 3 + r, 2 ** 40 + 12345 * r, 7 + r, 0]).cuda()
        states.append(st)
    want = sum(s.buffer.clone() for s in states)
    wires = [s.pack_wire().clone() for s in states]
    assert all(w.dtype == torch.int32 and w.numel() == 12 + C * C for w in wires)
    total = wires[0] + wires[1] + wires[2]
    assert int(total.min()) >= 0
    states[0].wire().copy_(total)
    states[0].unpack_wire()
    assert torch.equal(states[0].buffer, want)
    # rows counted on the host guard the exactness limit
    st = M.MetricState(C, k=5, device="cuda")
    st.update(torch.randn(10, C).cuda(), torch.zeros(10, dtype=torch.int64).cuda())
    st.note_rows(5)
    assert st.rows_fed == 15
    st.reset()
    assert st.rows_fed == 0
    assert N.lib().tb_metric_state_pack_i32(None, 4, None, None) == N.TB_EINVAL


@pytest.mark.parametrize("C,n", [(10, 1_000_000), (110, 200_000), (111, 200_000), (2, 300_000), (64, 4096)])
def test_confusion_bins_in_shared_memory_hot_bin_stress(C, n):
    """Few classes: the bins live in shared memory per CTA (C * C * 4 <= 48 KB, i.e. C <= 110; C = 111 is the first size
    on the global-atomic route).  A million samples over 100 bins, most of them on the diagonal, against bincount."""
    gen = torch.Generator(device="cuda").manual_seed(C)
    logits = torch.randn(n, C, device="cuda", generator=gen)
    labels = torch.randint(0, C, (n,), device="cuda", generator=gen)
    labels[::2] = logits[::2].argmax(-1)                       # half of the samples hit the diagonal
    labels[5::97] = -1                                          # a few invalid labels are skipped
    st = M.MetricState(C, k=min(3, C), device="cuda")
    st.update(logits, labels)
    st.update(logits[: n // 3], labels[: n // 3])               # accumulates across launches
    res = st.compute()
    L = torch.cat([logits, logits[: n // 3]]); Y = torch.cat([labels, labels[: n // 3]])
    ok = Y >= 0
    pred = L.argmax(-1)
    want = torch.bincount(Y[ok] * C + pred[ok], minlength=C * C).view(C, C)
    assert torch.equal(res["confusion"], want)
    assert res["correct_1"] == int((pred[ok] == Y[ok]).sum()) and res["invalid_labels"] == int((~ok).sum())
    assert res["n"] == n + n // 3


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16, torch.float16])
@pytest.mark.parametrize("C", [8, 64, 96, 1000, 1024])
def test_one_batch_rows_ties_nan_inf(C, dtype):
    """Rows the lean kernel takes (16-byte aligned, C a multiple of the vector width, C <= 1024) and the same values
    through the general kernel (a view whose rows start 16 bytes + one element off): heavy ties, NaN, +-inf, +-0,
    all-(-inf) rows, NaN label logits, labels at every position of a vector."""
    gen = torch.Generator().manual_seed(C)
    n = 257
    base = torch.randint(-3, 4, (n, C), generator=gen).float() / 2
    mask = torch.rand(n, C, generator=gen)
    base[mask < 0.01] = float("nan")
    base[(mask >= 0.01) & (mask < 0.02)] = float("inf")
    base[(mask >= 0.02) & (mask < 0.05)] = float("-inf")
    base[(mask >= 0.05) & (mask < 0.08)] = -0.0
    base[3] = float("-inf")                       # whole row -inf: argmax is index 0
    base[4] = 0.0; base[4, ::2] = -0.0            # zeros of both signs only
    base[5] = float("nan")                        # whole row NaN
    labels = torch.randint(0, C, (n,), generator=gen)
    labels[:C if C < n else n] = torch.arange(C)[:n]            # the label at every position of the first vectors
    labels[250:] = torch.tensor([-1, C, -5, C + 3, 0, C - 1, 1])
    x = base.to(dtype)
    ref = x.float()
    rank = MO.label_rank(ref, labels)
    valid = ((labels >= 0) & (labels < C)).numpy()
    wide = torch.zeros(n, C + 24, dtype=dtype)
    off = 8 // x.element_size() + 1                              # 8 bytes + one element: never 16-byte aligned
    wide[:, off:off + C] = x
    for logits in (x.cuda(), wide.cuda()[:, off:off + C]):
        st = M.MetricState(C, k=3, device="cuda")
        st.update(logits, labels.cuda())
        res = st.compute()
        assert res["correct_1"] == int(((rank < 1) & valid).sum())
        assert res["correct_3"] == int(((rank < 3) & valid).sum())
        assert res["invalid_labels"] == int((~valid).sum())
        assert np.array_equal(res["confusion"].cpu().numpy(), MO.confusion(ref, labels, C))
