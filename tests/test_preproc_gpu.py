"""GPU parity tests of the input side: keypoint heat maps against the golden vectors the
reference's own keypoint2heatmaps produced, and the Segment(20) path fed by them."""
import os

import numpy as np
import pytest
import torch

from instancesegmentation_b200 import Segment, synth
from instancesegmentation_b200.heatmaps import keypoint_heatmaps, stack_input
from oracle.heatmap_oracle import keypoint_heatmaps as oracle_heatmaps
from oracle.segment_oracle import segment_forward

from conftest import randomized_state

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _ulp_diff(a, b):
    ai, bi = a.view(np.int32).astype(np.int64), b.view(np.int32).astype(np.int64)
    return np.abs(ai - bi)


def test_heatmaps_match_reference_golden():
    """float64 exp on the device may differ from numpy's by one ulp of the double, i.e. virtually never after
    rounding to fp32: at most 1 fp32 ulp on at most 2 values per case, support identical elsewhere"""
    z = np.load(os.path.join(GOLDEN, "heatmaps_golden.npz"))
    for i in range(int(z["n"])):
        h, w, sigma, thr = z[f"meta{i}"]
        kp = torch.from_numpy(z[f"kp{i}"]).cuda()[None]
        got = keypoint_heatmaps(kp, (int(h), int(w)), sigma=float(sigma), threshold=float(thr))[0].cpu().numpy()
        want = z[f"maps{i}"]
        d = _ulp_diff(got, want)
        assert d.max() <= 1 and int((d != 0).sum()) <= 2, (i, int(d.max()), int((d != 0).sum()))
        assert int(((got != 0) != (want != 0)).sum()) <= 1, i


def test_heatmaps_batch_and_in_place_channel_slice():
    rng = np.random.default_rng(5)
    B, H, W = 5, 96, 128
    kp = np.zeros((B, 17, 3), np.float32)
    kp[..., 0] = rng.uniform(-20, W + 20, (B, 17))
    kp[..., 1] = rng.uniform(-20, H + 20, (B, 17))
    kp[..., 2] = rng.choice([0, 1, 1, 2], (B, 17))
    img = torch.from_numpy(synth.batch(B, H, W, seed=3)).cuda()
    x = stack_input(img, torch.from_numpy(kp).cuda())
    assert x.shape == (B, 20, H, W)
    assert torch.equal(x[:, :3], img)
    want = np.stack([oracle_heatmaps(kp[b], (H, W)) for b in range(B)])
    d = _ulp_diff(x[:, 3:].cpu().numpy(), want)
    assert d.max() <= 1 and int((d != 0).sum()) <= 4
    with pytest.raises(ValueError):
        keypoint_heatmaps(torch.zeros((1, 17, 2), device="cuda"), (H, W))


def test_segment20_train_batch_with_device_heatmaps():
    """Segment(20).train_batch(image, heatmaps) (segment.py:531-534) with the maps produced on the device"""
    ref = randomized_state(20, seed=5)
    sd = ref.state_dict()
    m = Segment(20)
    m.load_state_dict(sd, strict=True)
    m.precision, m.max_batch = "fp32", 2
    m = m.cuda().eval()
    B, H, W = 2, 64, 96
    rng = np.random.default_rng(8)
    kp = np.concatenate([rng.uniform(0, W, (B, 17, 1)), rng.uniform(0, H, (B, 17, 1)), np.ones((B, 17, 1))], 2).astype(np.float32)
    img = torch.from_numpy(synth.batch(B, H, W, seed=1))
    maps = keypoint_heatmaps(torch.from_numpy(kp).cuda(), (H, W))
    with torch.no_grad():
        got = m.train_batch(img.cuda(), maps).cpu()
    x = torch.cat([img, torch.from_numpy(np.stack([oracle_heatmaps(kp[b], (H, W)) for b in range(B)]))], 1)
    want = torch.sigmoid(segment_forward(sd, x))
    assert float((got - want).abs().max()) <= 1e-4


def test_crop_resize_bit_exact_against_oracle():
    from instancesegmentation_b200.preprocess import crop_resize, instance_windows
    from oracle import decode_oracle as D
    rng = np.random.default_rng(11)
    src = np.stack([synth.image_u8(96, 144, 40 + i) for i in range(3)])
    src[1] = rng.integers(0, 256, src[1].shape, dtype=np.uint8)           # white noise: every rounding case
    boxes = np.array([[10, 8, 90, 70], [0, 0, 144, 96], [-30, -20, 60, 50], [100, 60, 170, 120],
                      [40, 30, 47, 35], [300, 300, 340, 330], [5, 5, 6, 6]], np.int64)      # inside, full, over the edges, tiny, outside
    win = instance_windows(boxes, pad=16)
    inst_img = np.array([0, 1, 2, 0, 1, 2, 1], np.int32)
    for size in ((64, 64), (48, 80), (160, 208)):
        got = crop_resize(torch.from_numpy(src).cuda(), win, inst_img, size).cpu().numpy()
        want = D.crop_resize(src, win, inst_img, size)
        assert np.array_equal(got, want), (size, int((got != want).sum()))
    # a window of the network's own size is a plain copy with zero padding
    w1 = np.array([[-8, -4, 160, 112]], np.int32)
    got = crop_resize(torch.from_numpy(src).cuda(), w1, np.array([2], np.int32), (112, 160)).cpu().numpy()[0]
    assert np.array_equal(got[4:100, 8:152], src[2]) and got[:4].max() == 0 and got[:, :8].max() == 0 and got[100:].max() == 0
    with pytest.raises(ValueError):
        crop_resize(torch.from_numpy(src).cuda(), w1, np.array([5], np.int32), (112, 160))


def test_crop_forward_decode_round_trip():
    """infer.py chain on the device: windows -> crop_resize -> forward_u8 -> decode into the same windows;
    the pasted mask has the window's size and the statistics are in source coordinates"""
    from instancesegmentation_b200.decode import decode_masks
    from instancesegmentation_b200.preprocess import crop_resize, instance_windows
    z = np.load(os.path.join(GOLDEN, "segment3_state.npz"))
    m = Segment(3)
    m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}, strict=True)
    m.precision, m.max_batch = "fp32", 4
    m = m.cuda().eval()
    src = torch.from_numpy(np.stack([synth.image_u8(128, 160, 7), synth.image_u8(128, 160, 8)])).cuda()
    win = instance_windows(np.array([[20, 10, 100, 90], [60, 40, 150, 120], [0, 0, 50, 60]]), pad=16)
    inst_img = np.array([0, 0, 1], np.int32)
    crops = crop_resize(src, win, inst_img, (64, 96))
    with torch.no_grad():
        logits = m.forward_u8(crops, (0.5, 0.5, 0.5), (0.5, 0.5, 0.5))
        direct = m(((crops.permute(0, 3, 1, 2).float() / 255.0) - 0.5) / 0.5)
    assert float((logits - direct).abs().max()) <= 1e-4
    r = decode_masks(logits, win, img_first=np.array([0, 2, 3], np.int32), canvas=(128, 160))
    for n in range(3):
        assert tuple(r.mask(n).shape) == (int(win[n, 3]), int(win[n, 2]))
        if int(r.area[n]):
            bx = r.bbox[n].tolist()
            assert win[n, 0] <= bx[0] < bx[2] <= win[n, 0] + win[n, 2] and win[n, 1] <= bx[1] < bx[3] <= win[n, 1] + win[n, 3]
    assert r.idmap.shape == (2, 128, 160) and int(r.idmap.max()) <= 3


def test_infer_cli_end_to_end(tmp_path):
    """infer.py (reference flags, infer.py:12-29) over a directory of images of different sizes: every output is a
    0/255 PNG of its source's size and equals the library chain run by hand; --continue-test skips existing files"""
    import cv2
    import infer
    from instancesegmentation_b200.decode import decode_masks
    from instancesegmentation_b200.preprocess import crop_resize
    src_dir, out_dir = tmp_path / "in", tmp_path / "out"
    src_dir.mkdir()
    sizes = {"a.png": (70, 90), "b.jpg": (128, 64), "c.png": (33, 200)}
    for name, (h, w) in sizes.items():
        cv2.imwrite(str(src_dir / name), synth.image_u8(h, w, seed=h + w)[:, :, ::-1])
    ck = tmp_path / "ck.pth"
    z = np.load(os.path.join(GOLDEN, "segment3_state.npz"))
    sd = {k: torch.from_numpy(z[k]) for k in z.files}
    torch.save({"state_dict": sd}, ck)
    args = infer.parse_args(["-i", str(src_dir), "-o", str(out_dir), "--checkpoint", str(ck), "--size", "64", "--batch", "2"])
    assert infer.run(args) == 3
    m = Segment(3)
    m.load_state_dict(sd, strict=True)
    m.precision, m.max_batch = "fp32", 1
    m = m.cuda().eval()
    for name, (h, w) in sizes.items():
        got = cv2.imread(str(out_dir / (name[:-4] + ".png")), cv2.IMREAD_GRAYSCALE)
        assert got.shape == (h, w) and set(np.unique(got)) <= {0, 255}
        rgb = torch.from_numpy(np.ascontiguousarray(cv2.imread(str(src_dir / name), cv2.IMREAD_COLOR)[:, :, ::-1])).cuda()
        box = np.array([[0, 0, w, h]], np.int32)
        with torch.no_grad():
            r = decode_masks(m.forward_u8(crop_resize(rgb[None], box, np.zeros(1, np.int32), (64, 64)), (0.5,) * 3, (0.5,) * 3), box)
        assert np.array_equal(got, r.mask(0).cpu().numpy() * 255)
    args2 = infer.parse_args(["-i", str(src_dir), "-o", str(out_dir), "--checkpoint", str(ck), "--size", "64", "--continue-test"])
    assert infer.run(args2) == 0
