"""CPU tests: the oracle against the reference's own outputs (golden fixtures written by
tests/golden/make_golden.py from the unmodified reference, and the reference itself when
/root/reference is present)."""
import hashlib
import json
import os

import numpy as np
import torch
import torch.nn.functional as F

from conftest import GOLDEN
from instancesegmentation_b200 import synth
from oracle import decode_oracle as D
from oracle.segment_oracle import segment_forward

KEYS_SHA = "36e3e48511037b2aa7dfd8a2157f14fb05dcd0c6baa67b4c979f76567e94261b"


def test_golden_state_keys(golden_state):
    assert len(golden_state) == 588
    assert hashlib.sha256("\n".join(golden_state.keys()).encode()).hexdigest() == KEYS_SHA


def test_oracle_forward_matches_golden_logits(golden_state, golden_logits):
    meta = json.load(open(os.path.join(GOLDEN, "segment3_golden.json")))
    for name, c in meta["cases"].items():
        x = torch.from_numpy(synth.batch(c["B"], c["H"], c["W"], seed=c["seed"], kind=c["kind"]))
        y = segment_forward(golden_state, x).numpy()
        err = np.abs(y - golden_logits[name]).max()
        # same torch ops as the reference: bit-equal in the generating container; other CPUs
        # may pick different conv algorithms, hence the re-association allowance
        assert err <= 5e-5, (name, err)


def test_oracle_forward_matches_reference_module(reference_module, golden_state):
    m = reference_module.Segment(3)
    m.load_state_dict(golden_state, strict=True)
    m.eval()
    x = torch.from_numpy(synth.batch(2, 64, 112, seed=77))
    taps = {}
    with torch.no_grad():
        want = m(x)
    got = segment_forward(golden_state, x, taps)
    assert torch.equal(want, got)
    assert taps["init_down"].shape == (2, 19, 16, 28) and taps["bottle3_8"].shape == (2, 128, 4, 7)


def test_oracle_forward_in_channel_20(reference_module):
    torch.manual_seed(3)
    m = reference_module.Segment(20).eval()
    x = torch.randn(1, 20, 32, 48)
    with torch.no_grad():
        want = m(x)
    assert torch.equal(want, segment_forward(m.state_dict(), x))


def test_sigmoid_oracle_close_to_torch():
    rng = np.random.default_rng(0)
    x = np.concatenate([np.linspace(-30, 30, 6001), rng.normal(0, 2, 6000),
                        [0.0, 1e-30, -1e-30, 88.0, -88.0, 200.0, -200.0]]).astype(np.float32)
    s = D.sigmoid(x)
    t = torch.sigmoid(torch.from_numpy(x)).numpy()
    assert np.abs(s.astype(np.float64) - t).max() < 2.5e-7
    assert np.all(np.diff(s[:6001]) >= 0)
    assert s[-2] == 1.0 and s[-1] == 0.0
    assert np.isnan(D.sigmoid(np.array([np.nan], np.float32))[0])


def test_decode_oracle_against_torch_interpolate():
    L = synth.logits(4, 48, 64, seed=5)
    boxes = np.array([[0, 0, 64, 48], [3, 5, 128, 96], [1, 1, 23, 17], [7, 2, 100, 31]], np.int32)
    r = D.decode(L, boxes, want_soft=True)
    mism = 0
    for n, (x0, y0, wk, hk) in enumerate(boxes):
        p = F.interpolate(torch.sigmoid(torch.from_numpy(L[n:n + 1, None])), size=(int(hk), int(wk)),
                          mode="bilinear", align_corners=False)[0, 0].numpy()
        mine = r["masks"][r["off"][n]:r["off"][n + 1]].reshape(hk, wk)
        safe = np.abs(p - 0.5) > 1e-5           # torch differs from the restatement by <= ~2 ulp
        assert np.array_equal(mine[safe], (p > 0.5)[safe].astype(np.uint8))
        mism += int((mine != (p > 0.5)).sum())
        soft = r["soft"][r["off"][n]:r["off"][n + 1]].reshape(hk, wk).astype(np.int32)
        assert np.abs(soft - (p * 255).astype(np.int32)).max() <= 1
        assert r["area"][n] == int(mine.sum())
        ys, xs = np.nonzero(mine)
        if len(ys):
            assert list(r["bbox"][n]) == [x0 + xs.min(), y0 + ys.min(), x0 + xs.max() + 1, y0 + ys.max() + 1]
    assert mism <= 2


def test_decode_oracle_edge_cases():
    L = np.stack([np.full((16, 16), -5, np.float32), np.full((16, 16), 5, np.float32),
                  np.where(np.arange(256).reshape(16, 16) == 37, 4, -4).astype(np.float32)])
    boxes = np.array([[2, 3, 16, 16], [0, 0, 16, 16], [5, 5, 16, 16]], np.int32)
    r = D.decode(L, boxes)
    assert list(r["area"]) == [0, 256, 1]
    assert list(r["bbox"][0]) == [0, 0, 0, 0]
    assert list(r["bbox"][1]) == [0, 0, 16, 16]
    assert list(r["bbox"][2]) == [5 + 5, 5 + 2, 5 + 6, 5 + 3]
    # identity size: sigmoid(x) > 0.5 (not x > 0: tiny positive logits round to exactly 0.5)
    tiny = np.array([[[1e-9, 1e-6, -1e-6, 0.0]]], np.float32)
    assert list(D.decode(tiny, np.array([[0, 0, 4, 1]], np.int32))["masks"]) == [0, 1, 0, 0]


def test_idmap_oracle_rules():
    L = np.stack([np.full((8, 8), 1.0, np.float32), np.full((8, 8), 2.0, np.float32),
                  np.full((8, 8), 2.0, np.float32), np.full((8, 8), -1.0, np.float32)])
    boxes = np.array([[0, 0, 8, 8], [4, 0, 8, 8], [4, 4, 8, 8], [0, 0, 16, 16]], np.int32)
    ids = D.idmap(L, boxes, [0, 4], 12, 10)[0]
    assert ids[0, 0] == 1 and ids[0, 5] == 2      # higher probability wins
    assert ids[5, 5] == 2                          # tie between 2 and 3 -> lowest index
    assert ids[9, 9] == 3 and ids[11, 0] == 0      # box clipped by the canvas; unset mask -> 0
    assert D.mask_iou([1, 1, 0, 0], [1, 0, 1, 0]) == 1 / 3


def test_heatmap_oracle_matches_reference_function_golden():
    """oracle/heatmap_oracle.py against the vectors tests/golden/make_golden_heatmaps.py produced by
    executing the reference's own keypoint2heatmaps (train_instance.py:33-68): bit-exact."""
    from oracle.heatmap_oracle import keypoint_heatmaps
    z = np.load(os.path.join(GOLDEN, "heatmaps_golden.npz"))
    assert int(z["n"]) >= 4
    for i in range(int(z["n"])):
        h, w, sigma, thr = z[f"meta{i}"]
        got = keypoint_heatmaps(z[f"kp{i}"], (int(h), int(w)), sigma=int(sigma), threshold=float(thr))
        want = z[f"maps{i}"]
        assert got.dtype == np.float32 and got.shape == want.shape
        assert np.array_equal(got, want), i
        assert want[z[f"kp{i}"][:, 2] != 1].max(initial=0) == 0      # hidden / missing parts stay empty
        assert want[:, -1, :].max() == 0 and want[:, :, -1].max() == 0   # the reference never writes the last row/column


def test_bf16_tolerance_of_the_north_star_is_below_the_reference_graphs_own_bf16_sensitivity(golden_state):
    """north_star asks max-abs 2e-2 / mean-abs 2e-3 for bf16.  Evidence that no bf16-storage execution of
    this graph can meet it: the reference's fp32 arithmetic itself moves further than
    that when ONLY the input image, or ONLY the conv weights, are rounded to bf16 once.  The GPU test therefore
    asserts 'no worse than torch's own bf16 execution' and reports the numbers."""
    x = torch.from_numpy(synth.batch(1, 256, 256, seed=77))
    want = segment_forward(golden_state, x)
    d_in = (segment_forward(golden_state, x.to(torch.bfloat16).float()) - want).abs()
    sd_w = {k: (v.to(torch.bfloat16).float() if k.endswith("conv.weight") else v) for k, v in golden_state.items()}
    d_w = (segment_forward(sd_w, x) - want).abs()
    print(f"input->bf16: max {float(d_in.max()):.3f} mean {float(d_in.mean()):.4f} | "
          f"weights->bf16: max {float(d_w.max()):.3f} mean {float(d_w.mean()):.4f}")
    assert float(d_in.max()) > 2e-2 and float(d_in.mean()) > 2e-3
    assert float(d_w.max()) > 2e-2 and float(d_w.mean()) > 2e-3


def test_crop_resize_oracle_rules():
    """oracle_crop_resize (parity unpinned -- imgaug is absent): identity windows copy, everything outside the
    image is zero (CropAndPad), and the interpolation agrees with F.interpolate(bilinear, align_corners=False)
    on the zero-padded window to within the final rounding"""
    rng = np.random.default_rng(3)
    src = rng.integers(0, 256, (2, 40, 56, 3), dtype=np.uint8)
    same = D.crop_resize(src, np.array([[0, 0, 56, 40]], np.int32), np.array([1], np.int32), (40, 56))
    assert np.array_equal(same[0], src[1])
    out = D.crop_resize(src, np.array([[100, 100, 20, 20]], np.int32), np.array([0], np.int32), (16, 16))
    assert out.max() == 0
    win = np.array([[-7, -5, 50, 44]], np.int32)
    got = D.crop_resize(src, win, np.array([0], np.int32), (64, 80))[0].astype(np.float32)
    pad = np.zeros((44, 50, 3), np.float32)
    pad[5:44, 7:50] = src[0, 0:39, 0:43]
    want = F.interpolate(torch.from_numpy(pad).permute(2, 0, 1)[None], size=(64, 80), mode="bilinear",
                         align_corners=False)[0].permute(1, 2, 0).numpy()
    assert np.abs(got - want).max() <= 0.5 + 1e-3


def test_bce_oracle_matches_torch_bceloss_of_sigmoid():
    """pins oracle/bce_oracle.py on the reference's own calls (train_instance.py:299, :377-378): value and
    autograd gradient of BCELoss()(sigmoid(x), y) on CPU, including logits far enough out that the -100
    log clamp and the 1e-12 gradient clamp are in play"""
    from oracle.bce_oracle import bce_sigmoid
    rng = np.random.default_rng(5)
    x = (rng.standard_normal((2, 1, 37, 53)) * 4).astype(np.float32)
    x.reshape(-1)[:6] = [-200.0, 200.0, -90.0, 90.0, -20.0, 20.0]
    y = (rng.random(x.shape) > 0.7).astype(np.float32)
    y.reshape(-1)[:6] = [1, 0, 1, 0, 0, 1]
    for target in (y, rng.random(x.shape).astype(np.float32)):   # 0/1 masks, then soft targets
        xt = torch.from_numpy(x).requires_grad_(True)
        loss = torch.nn.BCELoss()(torch.sigmoid(xt), torch.from_numpy(target))
        loss.backward()
        o_loss, o_grad = bce_sigmoid(x, target)
        assert abs(float(o_loss) - loss.item()) <= 1e-6 * abs(loss.item())
        np.testing.assert_allclose(o_grad, xt.grad.numpy(), rtol=2e-5, atol=1e-10)
