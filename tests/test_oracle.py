"""The oracle against the golden vectors and against the third-party code it restates (CPU)."""
import numpy as np
import pytest
import torch
from PIL import Image

from oracle import metrics_oracle as MO
from oracle import params_oracle as PO
from oracle import resize_oracle as RO
from oracle import transform_oracle as TO

IMAGENET = ((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))
CIFAR = ((0.4914, 0.4822, 0.4465), (0.2023, 0.1994, 0.2010))


def bf16_from_bits(bits):
    return torch.from_numpy(np.ascontiguousarray(bits)).view(torch.bfloat16)


# ---- resize ---------------------------------------------------------------------------------
def test_resize_oracle_matches_golden(golden):
    g = golden("resize_pillow.npz")
    for i in range(int(g["n_cases"])):
        src, want = g[f"r{i}_in"], g[f"r{i}_out"]
        got = RO.resize_u8(src, want.shape[0], want.shape[1])
        assert np.array_equal(got, want), f"case {i}"


@pytest.mark.parametrize("shape,out", [((96, 128, 3), (48, 64)), ((75, 100, 3), (51, 68)), ((48, 64, 3), (52, 69)),
                                       ((40, 30), (13, 9))])
def test_resize_oracle_matches_pillow_live(shape, out):
    rng = np.random.default_rng(5)
    img = rng.integers(0, 256, shape, dtype=np.uint8)
    want = np.asarray(Image.fromarray(img).resize((out[1], out[0]), Image.BILINEAR))
    assert np.array_equal(RO.resize_u8(img, *out), want)


def test_taps_512_to_256_known_answer():
    # SURVEY.md section 8(a) T5
    ks, bounds, kk = RO.bilinear_taps(512, 256)
    assert ks == 5
    assert kk[100, :4].tolist() == [524288, 1572864, 1572864, 524288]
    assert kk[0, :3].tolist() == [1797559, 1797559, 599186]
    assert bounds[0].tolist() == [0, 3]


def test_resized_output_size():
    assert RO.resized_output_size(375, 500, 256) == (256, 341)
    assert RO.resized_output_size(480, 640, 512) == (512, 682)
    assert RO.resized_output_size(512, 512, 256) == (256, 256)
    assert RO.resized_output_size(300, 200, (384, 256)) == (384, 256)


# ---- transform ------------------------------------------------------------------------------
def test_transform_oracle_matches_golden(golden):
    g = golden("transform_torchvision.npz")
    out = TO.transform_batch(g["c1_in"], crop_top=g["c1_top"], crop_left=g["c1_left"], out_size=(32, 32),
                             flip=g["c1_flip"], padding=4, padding_mode="reflect", mean=CIFAR[0], std=CIFAR[1])
    assert np.array_equal(out.numpy(), g["c1_out"])
    out = TO.transform_batch(g["c2_in"], crop_top=g["c2_top"], crop_left=g["c2_left"], out_size=(40, 56),
                             flip=g["c2_flip"], mean=IMAGENET[0], std=IMAGENET[1], dtype=torch.bfloat16)
    assert torch.equal(out, bf16_from_bits(g["c2_out_bf16_bits"]))
    out = TO.transform_batch(g["c3_in"], flip=g["c3_flip"], mean=[.5] * 3, std=[.5] * 3)
    assert np.array_equal(out.numpy(), g["c3_out"])
    out = TO.transform_batch(g["gan_in"], one_minus=True)
    assert np.array_equal(out.numpy(), g["gan_out"])
    out = TO.transform_batch(g["c4_in"], resize=32, center_crop=32, mean=IMAGENET[0], std=IMAGENET[1])
    assert np.array_equal(out.numpy(), g["c4_out"])


@pytest.mark.parametrize("mode", ["constant", "reflect", "edge", "symmetric"])
def test_index_map_equals_torchvision(mode):
    """The fused index map + 256-entry table is the torchvision chain, exactly."""
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (20, 24, 3), dtype=np.uint8)
    lut = np.stack([((np.arange(256, dtype=np.float32) / np.float32(255)) - np.float32(m)) / np.float32(s)
                    for m, s in zip(*IMAGENET)]).astype(np.float32)
    for top, left, flip in [(0, 0, False), (6, 10, True), (3, 0, True), (6, 5, False)]:
        want = TO.transform_sample(img, padding=3, padding_mode=mode, crop=(top, left, 20, 20), flip=flip,
                                   mean=IMAGENET[0], std=IMAGENET[1])
        got = TO.index_map_sample(img, lut, top=top, left=left, out_h=20, out_w=20, flip=flip, pad_y=3, pad_x=3,
                                  pad_mode=mode)
        assert np.array_equal(got, want.numpy())


def test_fused_scale_bias_is_not_parity_safe():
    """x*scale+bias differs from torchvision's (x/255 - mean)/std -- hence the table."""
    x = np.arange(256, dtype=np.float32)
    m, s = np.float32(0.485), np.float32(0.229)
    ref = (x / np.float32(255) - m) / s
    fused = x * (np.float32(1) / (np.float32(255) * s)) + (-m / s)
    assert (ref != fused).sum() > 50


# ---- params ---------------------------------------------------------------------------------
def test_params_oracle_matches_golden(golden):
    g = golden("params_torchvision.npz")
    for name in ("c2", "c1", "eq", "rect"):
        seed, B, h, w, th, tw, pad = g[f"{name}_cfg"].tolist()
        top, left, flip = PO.draw_crop_flip(seed, B, h, w, th, tw, pad)
        want = g[f"{name}_params"]
        assert np.array_equal(top, want[:, 0]) and np.array_equal(left, want[:, 1]) and np.array_equal(flip, want[:, 2])
    assert g["c2_params"][:3].tolist() == [[6, 29, 1], [25, 24, 0], [25, 11, 0]]


# ---- metrics --------------------------------------------------------------------------------
def test_metrics_oracle_matches_reference_golden(golden):
    g = golden("metrics_reference.npz")
    kat = torch.from_numpy(g["kat_logits"])
    assert np.array_equal(MO.argmax_first(kat), g["kat_argmax"])
    assert g["kat_argmax"][:4].tolist() == [1, 0, 0, 2]
    for lab in range(4):
        labels = torch.full((kat.shape[0],), lab, dtype=torch.int64)
        assert MO.accuracy(kat, labels).numpy() == g[f"kat_accuracy_label{lab}"]
        # rank rule with k=1 is the reference's argmax rule
        assert MO.topk_correct(kat, labels, 1) == round(float(g[f"kat_accuracy_label{lab}"]) * kat.shape[0])
    a = MO.accuracy(torch.from_numpy(g["a_logits"]), torch.from_numpy(g["a_labels"]))
    assert a.numpy() == g["a_accuracy"] and a.dtype == torch.float32
    lt = bf16_from_bits(g["b_logits_bf16_bits"])
    lb = torch.from_numpy(g["b_labels"])
    assert MO.accuracy(lt, lb).float().numpy() == g["b_accuracy"]
    assert MO.topk_correct(lt, lb, 1) == round(float(g["b_accuracy"]) * 300)
    assert MO.accuracy(torch.from_numpy(g["c_logits"]), torch.from_numpy(g["c_labels"])).numpy() == g["c_accuracy"] == 5.0
    ra = MO.RunningAverage()
    for v, want in zip(g["ra_inputs"], g["ra_values"]):
        ra.update(float(v))
        assert ra.value == want
    assert g["ra_123"].tolist() == [0.20000000000000004, 3.0]


def test_rank_rule_agrees_with_topk_and_bincount_without_ties():
    gen = torch.Generator().manual_seed(0)
    logits = torch.randn(500, 37, generator=gen)
    labels = torch.randint(0, 37, (500,), generator=gen)
    for k in (1, 3, 5):
        want = (logits.topk(k, dim=-1).indices == labels[:, None]).any(-1).sum().item()
        assert MO.topk_correct(logits, labels, k) == want
    pred = logits.argmax(-1)
    want = torch.bincount(labels * 37 + pred, minlength=37 * 37).view(37, 37)
    assert np.array_equal(MO.confusion(logits, labels, 37), want.numpy())
    # SURVEY.md 8(c) 4-sample example
    lg = torch.tensor([[9., 0, 0], [0, 0, 9], [0, 0, 9], [0, 9, 0]])
    assert MO.confusion(lg, torch.tensor([0, 1, 2, 2]), 3).tolist() == [[1, 0, 0], [0, 0, 1], [0, 1, 1]]


def test_rank_rule_ties_and_nan():
    nan = float("nan")
    lg = torch.tensor([[2., 2, 2, 2], [nan, 5, 1, nan], [1, nan, 0, 0]])
    assert MO.label_rank(lg, torch.tensor([2, 3, 0])).tolist() == [2, 1, 1]
    assert MO.label_rank(lg, torch.tensor([0, 0, 1])).tolist() == [0, 0, 0]
    assert MO.label_rank(lg, torch.tensor([7, -1, 0])).tolist() == [4, 4, 1]


def test_resized_crop_oracle_matches_golden_and_integer_restatement(golden):
    g = golden("transform_torchvision.npz")
    out = TO.transform_batch(g["rrc_in"], boxes=g["rrc_boxes"], out_size=(24, 24), flip=g["rrc_flip"],
                             mean=IMAGENET[0], std=IMAGENET[1])
    assert np.array_equal(out.numpy(), g["rrc_out"])
    # F.resized_crop == integer resample of the cropped box
    for b in range(3):
        i, j, h, w = g["rrc_boxes"][b]
        rs = RO.resize_u8(np.ascontiguousarray(g["rrc_in"][b][i:i + h, j:j + w]), 24, 24)
        if g["rrc_flip"][b]:
            rs = rs[:, ::-1]
        ref = torch.from_numpy(np.ascontiguousarray(rs)).permute(2, 0, 1).float().div(255)
        ref = (ref - torch.tensor(IMAGENET[0]).view(3, 1, 1)) / torch.tensor(IMAGENET[1]).view(3, 1, 1)
        assert torch.equal(ref, out[b])
