"""GPU parity tests of the forward: CUDA engine (through the C ABI) vs the oracle restatement
and the committed reference logits.  Tolerances: fp32 mode 1e-4 max-abs on logits
(BASELINE.json north_star); tensor-core mode "bf16x3" (split-precision operands) max-abs 2e-2 /
mean-abs 2e-3 (north_star's bf16 tolerance); plain "bf16" is an extra mode OUTSIDE that tolerance
(reported, sanity-bounded), see DESIGN.md section 5."""
import json
import os

import numpy as np
import pytest
import torch

from conftest import GOLDEN, randomized_state
from instancesegmentation_b200 import Segment, synth
from oracle.segment_oracle import segment_forward

pytestmark = pytest.mark.gpu
FP32_TOL = 1e-4


def _fp32_check(sd, x, got, label=""):
    """fp32-mode acceptance.  The reference's own fp32 arithmetic sits
    1.2e-4 (max-abs, 512x512) from the float64 value of its graph and 6e-5 from itself when the
    thread count changes, so "within 1e-4 of the reference" is only decidable up to that noise:
    we require (a) max-abs <= 1e-4 against the fp32 oracle OR our distance to the float64
    evaluation to be no larger than 1.25x the oracle's own distance to it, and always (b) a hard
    cap of 2.5e-4 and mean-abs <= 1.5e-5 against the fp32 oracle."""
    want32 = segment_forward(sd, x)
    d32 = (got - want32).abs()
    msg = f"{label} vs fp32 oracle: max {float(d32.max()):.3e} mean {float(d32.mean()):.3e}"
    ok = float(d32.max()) <= FP32_TOL
    if not ok:
        want64 = segment_forward(sd, x, dtype=torch.float64)
        ours64 = float((got.double() - want64).abs().max())
        ref64 = float((want32.double() - want64).abs().max())
        msg += f" | vs float64: ours {ours64:.3e}, fp32 oracle itself {ref64:.3e}"
        ok = ours64 <= 1.25 * ref64
    print(msg)
    assert ok, msg
    assert float(d32.max()) <= 2.5e-4 and float(d32.mean()) <= 1.5e-5, msg


def _model(sd, cin=3, precision="fp32", max_batch=8):
    m = Segment(cin)
    m.load_state_dict(sd, strict=True)
    m.precision, m.max_batch = precision, max_batch
    return m.cuda().eval()


def _engine(m):
    return next(iter(m._engines.values()))


def test_fp32_matches_committed_reference_logits(golden_state, golden_logits):
    meta = json.load(open(os.path.join(GOLDEN, "segment3_golden.json")))
    m = _model(golden_state)
    for name, c in meta["cases"].items():
        x = torch.from_numpy(synth.batch(c["B"], c["H"], c["W"], seed=c["seed"], kind=c["kind"])).cuda()
        with torch.no_grad():
            y = m(x)
        assert y.shape == (c["B"], 1, c["H"], c["W"]) and y.dtype == torch.float32 and y.is_cuda
        err = np.abs(y.cpu().numpy() - golden_logits[name]).max()
        scale = max(1.0, c["logit_absmax"] / 3.0)   # the noise case has logits ~4x larger
        print(f"{name}: max-abs err {err:.3e}")
        assert err <= FP32_TOL * scale, (name, err)


def test_fp32_layerwise_against_oracle(golden_state):
    m = _model(golden_state)
    x = torch.from_numpy(synth.batch(2, 96, 160, seed=4321))
    taps = {}
    want = segment_forward(golden_state, x, taps)
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    mine = _engine(m).taps(2)
    report = {}
    for k, t in mine.items():
        assert k in taps, k
        assert t.shape == taps[k].shape, (k, t.shape, taps[k].shape)
        report[k] = float((t - taps[k]).abs().max())
    print({k: f"{v:.2e}" for k, v in report.items()})
    first_bad = next((k for k, v in report.items() if v > 1e-4 * max(1.0, float(taps[k].abs().max()))), None)
    assert first_bad is None, (first_bad, report)
    assert float((got - want).abs().max()) <= FP32_TOL


@pytest.mark.parametrize("B,H,W", [(1, 16, 16), (3, 64, 64), (1, 480, 480), (2, 512, 512), (5, 32, 208)])
def test_fp32_sizes_and_batches(golden_state, B, H, W):
    m = _model(golden_state, max_batch=2)   # B=3,5 exercise the max_batch grouping
    x = torch.from_numpy(synth.batch(B, H, W, seed=9000 + H))
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    _fp32_check(golden_state, x, got, f"B{B} {H}x{W}")


def test_fp32_1024(golden_state):
    m = _model(golden_state, max_batch=1)
    x = torch.from_numpy(synth.batch(1, 1024, 1024, seed=31))
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    _fp32_check(golden_state, x, got, "1x1024x1024")


def test_fp32_in_channel_20():
    ref = randomized_state(20, seed=5)
    sd = ref.state_dict()
    m = _model(sd, cin=20)
    x = torch.from_numpy(synth.batch(2, 64, 96, seed=1, channels=20))
    want = segment_forward(sd, x)
    with torch.no_grad():
        got = m.train_batch(x[:, :3].cuda(), x[:, 3:].cuda()).cpu()   # segment.py:531-534
    assert float((got - torch.sigmoid(want)).abs().max()) <= 2e-5
    with torch.no_grad():
        assert float((m(x.cuda()).cpu() - want).abs().max()) <= FP32_TOL


def test_weight_reload_and_batch_independence(golden_state):
    m = _model(golden_state)
    x = torch.from_numpy(synth.batch(4, 64, 64, seed=50)).cuda()
    with torch.no_grad():
        y4 = m(x)
        y1 = torch.cat([m(x[i:i + 1]) for i in range(4)])
        assert torch.equal(y4, y1)          # batch sharding is exact
        sd2 = {k: (v * 1.01 if v.dtype.is_floating_point and k.endswith("conv.weight") else v)
               for k, v in golden_state.items()}
        m.load_state_dict(sd2)
        y2 = m(x)
    assert float((y2 - y4).abs().max()) > 1e-4
    assert float((y2.cpu() - segment_forward(sd2, x.cpu())).abs().max()) <= FP32_TOL


@pytest.mark.parametrize("H,W", [(96, 160), (512, 512), (208, 48), (16, 16), (1024, 1024), (64, 2048), (1536, 32)])
def test_bf16_layerwise_against_oracle(golden_state, H, W):
    """Every tapped activation of the tensor-core path stays within bf16 noise of the oracle
    (a wrong descriptor, tap offset or weight slot shows up as an O(1) relative error); sizes from the
    smallest legal one to cfg4's 1024x1024 and very wide / very tall images."""
    m = _model(golden_state, precision="bf16", max_batch=2)
    x = torch.from_numpy(synth.batch(2, H, W, seed=4321))
    taps = {}
    segment_forward(golden_state, x, taps)
    with torch.no_grad():
        m(x.cuda())
    mine = _engine(m).taps(2)
    report = {k: float((t - taps[k]).abs().max() / taps[k].abs().max().clamp_min(1e-6)) for k, t in mine.items()}
    print({k: f"{v:.1e}" for k, v in report.items()})
    bad = {k: v for k, v in report.items() if not v < 0.3}
    assert not bad, bad


X3_MAX, X3_MEAN = 2e-2, 2e-3    # north_star: logits within max-abs 2e-2 / mean-abs 2e-3 on the tensor-core path


@pytest.mark.parametrize("B,H,W", [(2, 512, 512), (1, 1024, 1024), (2, 96, 160), (3, 64, 64), (1, 16, 16)])
def test_bf16x3_logits_within_north_star_tolerance(golden_state, B, H, W):
    """The tensor-core mode the bench measures: split-precision bf16 operands (hi + lo, three MMA passes)."""
    x = torch.from_numpy(synth.batch(B, H, W, seed=77 + H))
    want = segment_forward(golden_state, x)
    m = _model(golden_state, precision="bf16x3", max_batch=2)
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    d = (got - want).abs()
    print(f"bf16x3 {B}x{H}x{W}: max {float(d.max()):.3e} mean {float(d.mean()):.3e}")
    assert float(d.max()) <= X3_MAX and float(d.mean()) <= X3_MEAN


def test_bf16x3_matches_committed_reference_logits(golden_state, golden_logits):
    """the same bar against logits written by the unmodified reference module (tests/golden/make_golden.py)"""
    meta = json.load(open(os.path.join(GOLDEN, "segment3_golden.json")))
    m = _model(golden_state, precision="bf16x3")
    for name, c in meta["cases"].items():
        x = torch.from_numpy(synth.batch(c["B"], c["H"], c["W"], seed=c["seed"], kind=c["kind"])).cuda()
        with torch.no_grad():
            y = m(x)
        d = np.abs(y.cpu().numpy() - golden_logits[name])
        scale = max(1.0, c["logit_absmax"] / 3.0)   # the noise case has logits ~4x larger
        print(f"{name}: max-abs {d.max():.3e} mean-abs {d.mean():.3e}")
        assert d.max() <= X3_MAX * scale and d.mean() <= X3_MEAN * scale, name


@pytest.mark.parametrize("H,W", [(96, 160), (512, 512), (208, 48), (16, 16), (1024, 1024), (64, 2048), (1536, 32)])
def test_bf16x3_layerwise_against_oracle(golden_state, H, W):
    """Every tapped activation within 2e-3 of the tensor's scale (measured <= 3.2e-4): a wrong descriptor, tap
    offset, weight slot or hi/lo plane shows up orders of magnitude above that, also in a low-energy channel."""
    m = _model(golden_state, precision="bf16x3", max_batch=2)
    x = torch.from_numpy(synth.batch(2, H, W, seed=4321))
    taps = {}
    segment_forward(golden_state, x, taps)
    with torch.no_grad():
        m(x.cuda())
    mine = _engine(m).taps(2)
    report = {k: float((t - taps[k]).abs().max() / taps[k].abs().max().clamp_min(1e-6)) for k, t in mine.items()}
    print({k: f"{v:.1e}" for k, v in report.items()})
    bad = {k: v for k, v in report.items() if not v < 2e-3}
    assert not bad, bad


def test_plain_bf16_mode_is_labelled_out_of_tolerance(golden_state):
    """Plain bf16 operands (ISG_MODE_BF16) do NOT meet the north_star tolerance on this network (rounding
    only the input image to bf16 already moves the logits by 0.25); the mode is kept as a faster, labelled extra.
    This test reports its error beside torch's own bf16 execution and only sanity-bounds it."""
    x = torch.from_numpy(synth.batch(2, 256, 256, seed=77))
    want = segment_forward(golden_state, x)
    m = _model(golden_state, precision="bf16")
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    ours = (got - want).abs()
    ref = Segment(3)
    ref.load_state_dict(golden_state)
    ref = ref.cuda().train(False)
    with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
        tb = ref._forward_autograd(x.cuda()).float().cpu()
    theirs = (tb - want).abs()
    print(f"plain bf16 (out of tolerance) max {ours.max():.3e} mean {ours.mean():.3e} | torch autocast max "
          f"{theirs.max():.3e} mean {theirs.mean():.3e}")
    assert float(ours.mean()) <= 0.1 and float(ours.max()) <= 1.5     # sanity only; the contract is bf16x3's
    assert float(ours.max()) > X3_MAX                                   # if this ever fails, drop the label


def _u8_batch(B, H, W, seed):
    rng = np.random.default_rng(seed)
    return torch.from_numpy(rng.integers(0, 256, size=(B, H, W, 3), dtype=np.uint8))


def _normalize_like_reference(u8, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5)):
    """ToTensor + Normalize of train_instance.py:80-85 on the host in fp32."""
    x = u8.permute(0, 3, 1, 2).float().div(255.0)
    m = torch.tensor(mean).view(1, 3, 1, 1)
    s = torch.tensor(std).view(1, 3, 1, 1)
    return x.sub(m).div(s).contiguous()


@pytest.mark.parametrize("precision", ["fp32", "bf16", "bf16x3"])
def test_forward_u8_equals_forward_of_normalized_input(golden_state, precision):
    """isg_forward_u8 (uint8 NHWC in, Normalize on the device) must give exactly the logits of
    isg_forward on the host-normalised fp32 tensor: the normalisation is the same IEEE sequence."""
    u8 = _u8_batch(3, 64, 96, seed=5)
    x = _normalize_like_reference(u8, (0.4, 0.5, 0.6), (0.5, 0.25, 0.2))
    m = _model(golden_state, precision=precision)
    with torch.no_grad():
        want = m(x.cuda()).cpu()
        got = m.forward_u8(u8.cuda(), (0.4, 0.5, 0.6), (0.5, 0.25, 0.2)).cpu()
    assert torch.equal(got, want), float((got - want).abs().max())
    if precision == "fp32":
        _fp32_check(golden_state, x, got, "u8 64x96")


_MODE_SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1])
from instancesegmentation_b200 import Segment, synth
z = np.load(sys.argv[2])
B, H, W = (int(v) for v in sys.argv[4].split("x"))
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = sys.argv[5], B
m = m.cuda().eval()
x = torch.from_numpy(synth.batch(B, H, W, seed=3)).cuda()
with torch.no_grad():
    y = m(x)
torch.cuda.synchronize()
np.save(sys.argv[3], y.cpu().numpy())
"""


@pytest.mark.parametrize("size,envs", [
    ("4x128x192", [{"ISG_FK_NO_TUNE": "1"}, {}, {"ISG_FK_CTAS": "2", "ISG_FK_LBUF": "1"},
                   {"ISG_FK_CTAS": "2", "ISG_FK_LBUF": "2"}, {"ISG_FK_CTAS": "1", "ISG_FK_LBUF": "1"},
                   {"ISG_FK_FORCE_PIPE": "1"}, {"ISG_FK_CTAS": "4", "ISG_FK_LBUF": "2"},
                   {"ISG_FK_CTAS": "4", "ISG_FK_LBUF": "1"}, {"ISG_FK_NO_PLANAR": "1"}]),
    # 512x512: the residual-bottleneck runs go through the cluster-resident launch (csrc/mega.cuh)
    # ISG_FK_NO_PLANAR: pixel-major instead of planar global tensors (only the TMA box shape differs)
    ("3x512x512", [{"ISG_FK_NO_MEGA": "1", "ISG_FK_NO_TUNE": "1"}, {}, {"ISG_FK_NO_MEGA8": "1"}, {"ISG_FK_NO_PLANAR": "1"}]),
    # 1024x1024 (cfg4): the untuned default mode has no feasible tile for one launch -> fallback search
    ("2x1024x1024", [{"ISG_FK_NO_TUNE": "1"}, {}]),
    # split-precision instantiations of the same kernels
    ("bf16x3:4x128x192", [{"ISG_FK_NO_TUNE": "1"}, {}, {"ISG_FK_CTAS": "2", "ISG_FK_LBUF": "1"}, {"ISG_FK_CTAS": "1", "ISG_FK_LBUF": "1"},
                          {"ISG_FK_FORCE_PIPE": "1"}, {"ISG_FK_CTAS": "4", "ISG_FK_LBUF": "2"}, {"ISG_FK_NO_PLANAR": "1"}]),
    # 512x512: the residual runs go through the 8-CTA-cluster launch (csrc/mega3.cuh) unless ISG_FK_NO_MEGA
    # (per-block launches hand m1 over as a hi + lo pair, mega3 keeps it in fp32: equal within 5e-3 (measured 1.1e-3), not bitwise)
    ("bf16x3:3x512x512", [{"ISG_FK_NO_TUNE": "1"}, {}, {"ISG_FK_NO_MEGA": "1", "ISG_FK_NO_TUNE": "1", "TOL": "5e-3"},
                          {"ISG_FK_NO_MEGA8": "1", "TOL": "5e-3"}]),
])
def test_bf16_launch_variants_are_bitwise_identical(tmp_path, size, envs):
    """Every way a fused launch may run (tile, CTAs per SM, load buffering -- the autotuner's choices -- and
    per-block launches vs one cluster launch per bottleneck run) computes each pixel with the same
    instruction sequence: logits must not depend on the choice."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "run.py"
    script.write_text(_MODE_SCRIPT)
    outs = []
    precision, size = size.split(":") if ":" in size else ("bf16", size)
    for i, env in enumerate(envs):
        out = tmp_path / f"y{i}.npy"
        e = dict(os.environ, **env)
        subprocess.run([sys.executable, str(script), root, os.path.join(GOLDEN, "segment3_state.npz"), str(out), size, precision],
                       check=True, env=e, timeout=600)
        outs.append(np.load(out))
    for i in range(1, len(outs)):
        d = float(np.abs(outs[0] - outs[i]).max())
        if "TOL" in envs[i]:
            assert d <= float(envs[i]["TOL"]), (i, d)
        else:
            assert np.array_equal(outs[0], outs[i]), (i, d)


def test_pipeline_matches_direct_calls(golden_state):
    """Three-stream serving pipeline (fp32 and uint8 batches): same masks / stats as plain calls."""
    from instancesegmentation_b200.decode import decode_masks
    from instancesegmentation_b200.pipeline import Pipeline, identity_boxes
    B, H, W = 4, 64, 64
    m = _model(golden_state, precision="bf16", max_batch=B)
    pipe = Pipeline(m, identity_boxes(B, H, W), depth=2)
    batches = [_u8_batch(B, H, W, seed=s).pin_memory() for s in range(3)]
    batches += [_normalize_like_reference(b).pin_memory() for b in batches[:2]]
    outs = [(torch.empty(B * H * W, dtype=torch.uint8).pin_memory(), torch.empty(B, dtype=torch.int32).pin_memory(),
             torch.empty((B, 4), dtype=torch.int32).pin_memory()) for _ in batches]
    for hb, (mh, ah, bh) in zip(batches, outs):
        pipe.submit(hb, mh, (ah, bh))
    pipe.finish()
    torch.cuda.synchronize()
    for hb, (mh, ah, bh) in zip(batches, outs):
        with torch.no_grad():
            logits = m.forward_u8(hb.cuda()) if hb.dtype == torch.uint8 else m(hb.cuda())
            r = decode_masks(logits)
        assert torch.equal(mh, r.masks.cpu()[:B * H * W])
        assert torch.equal(ah, r.area.cpu()) and torch.equal(bh, r.bbox.cpu())
    # uint8 and host-normalised fp32 batches of the same images agree
    assert torch.equal(outs[0][0], outs[3][0]) and torch.equal(outs[1][0], outs[4][0])
    # graph replay + bit-packed masks: same result, 8x fewer bytes back
    pipe2 = Pipeline(m, identity_boxes(B, H, W), depth=2, graphs=True, packed=True)
    nb = (B * H * W + 7) // 8
    packed = [torch.empty(nb, dtype=torch.uint8).pin_memory() for _ in range(3)]
    for hb, ph in zip(batches[:3], packed):
        pipe2.submit(hb, ph)
    pipe2.finish()
    torch.cuda.synchronize()
    for (mh, _, _), ph in zip(outs[:3], packed):
        assert np.array_equal(np.unpackbits(ph.numpy(), bitorder="little")[:B * H * W], mh.numpy())


_BLOCK_TAP_SCRIPT = r"""
import sys, numpy as np, torch
sys.path.insert(0, sys.argv[1])
from instancesegmentation_b200 import Segment, synth
from oracle.segment_oracle import segment_forward
z = np.load(sys.argv[2])
sd = {k: torch.from_numpy(z[k]) for k in z.files}
prec, H, W = sys.argv[3], int(sys.argv[4]), int(sys.argv[5])
m = Segment(3); m.load_state_dict(sd); m.precision, m.max_batch = prec, 2
m = m.cuda().eval()
x = torch.from_numpy(synth.batch(2, H, W, seed=4321))
taps = {}
segment_forward(sd, x, taps)
with torch.no_grad():
    m(x.cuda())
mine = next(iter(m._engines.values())).taps(2)
blocks = [k for k in mine if ".x." in k.replace("_x.", ".x.")]
assert len(blocks) >= 11, sorted(mine)
for k, t in mine.items():
    print(k, float((t - taps[k]).abs().max() / taps[k].abs().max().clamp_min(1e-6)))
"""


@pytest.mark.parametrize("precision,bound", [("bf16x3", 2e-3), ("bf16", 0.3)])
def test_every_residual_block_is_tapped_with_per_block_launches(tmp_path, precision, bound):
    """With ISG_FK_NO_MEGA=1 every Bottleneck3x3 / Bottleneck5x5 (segment.py:52-111) is its own launch and its output a
    tap: each of the 14 blocks is compared with the oracle on its own (the Bottleneck5x5 blocks bottle2_x.4 = bottle2_8 and
    bottle3_x.4 = bottle3_8 with their position-dependent bias included).  The cluster-resident launches are bitwise
    identical to these (test_bf16_launch_variants_are_bitwise_identical)."""
    import subprocess, sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    script = tmp_path / "taps.py"
    script.write_text(_BLOCK_TAP_SCRIPT)
    env = dict(os.environ, ISG_FK_NO_MEGA="1", ISG_FK_NO_TUNE="1")
    out = subprocess.run([sys.executable, str(script), root, os.path.join(GOLDEN, "segment3_state.npz"), precision, "512", "512"],
                         check=True, env=env, timeout=600, capture_output=True, text=True).stdout
    report = {l.split()[0]: float(l.split()[1]) for l in out.splitlines() if len(l.split()) == 2}
    print(report)
    for k in ["bottle1_x.0", "bottle1_x.2", "bottle2_x.0", "bottle2_x.1", "bottle2_x.3", "bottle3_x.1", "bottle3_x.3",
              "bottle2_8", "bottle3_8"]:
        assert k in report, (k, sorted(report))
    bad = {k: v for k, v in report.items() if not v < bound}
    assert not bad, bad


def test_one_plan_per_size_serves_smaller_batches(golden_state):
    """The engine cache is keyed by (device, size, precision): a smaller batch reuses the plan (no second autotune /
    workspace), a larger one replaces it; results do not depend on which plan served them."""
    m = _model(golden_state, precision="bf16x3", max_batch=8)
    x = torch.from_numpy(synth.batch(6, 64, 64, seed=21)).cuda()
    with torch.no_grad():
        y3 = m(x[:3])
        e3 = _engine(m)
        y1 = m(x[:1])
        assert _engine(m) is e3 and len(m._engines) == 1          # B = 1 rides on the plan built for 3
        y6 = m(x)
        e6 = _engine(m)
        assert e6 is not e3 and e6.max_batch == 6 and len(m._engines) == 1
        assert torch.equal(m(x[:3]), y3) and _engine(m) is e6
    assert torch.equal(y1, y3[:1]) and torch.equal(y6[:3], y3)


def test_bf16x3_exact_uint8_first_layer(golden_state):
    """isg_forward_u8 with the reference's Normalize(0.5, 0.5) (train_instance.py:80-85) on a split-precision plan takes
    the exact first layer: centred pixels v - 127.5 as single bf16 values, 1 / (255 std) folded into layer1's weights.
    Same bar against the oracle on the host-normalised tensor; agrees with the generic path (other mean / std fall back
    to it, and forward() of the normalised fp32 tensor always uses it) to rounding; re-packs when mean / std change."""
    u8 = _u8_batch(2, 96, 160, seed=8)
    x = _normalize_like_reference(u8)
    want = segment_forward(golden_state, x)
    m = _model(golden_state, precision="bf16x3")
    with torch.no_grad():
        exact = m.forward_u8(u8.cuda()).cpu()                       # 0.5 / 0.5 -> exact path
        generic = m(x.cuda()).cpu()
        d = (exact - want).abs()
        print(f"exact-u8 path: max {float(d.max()):.3e} mean {float(d.mean()):.3e}; vs generic {float((exact - generic).abs().max()):.3e}")
        assert float(d.max()) <= X3_MAX and float(d.mean()) <= X3_MEAN
        assert float((exact - generic).abs().max()) <= 2e-3
        # integer centre (mean = 128/255) is exact too, with another std: the variant is re-packed
        mean2, std2 = (128.0 / 255.0,) * 3, (0.25, 0.5, 0.2)
        x2 = _normalize_like_reference(u8, mean2, std2)
        got2 = m.forward_u8(u8.cuda(), mean2, std2).cpu()
        d2 = (got2 - segment_forward(golden_state, x2)).abs()
        assert float(d2.max()) <= X3_MAX * 2 and float(d2.mean()) <= X3_MEAN * 2     # inputs up to 4x larger
        again = m.forward_u8(u8.cuda()).cpu()                       # back to 0.5 / 0.5: packed again, same result
        assert torch.equal(again, exact)
        # a mean whose centred pixels are not bf16 values falls back to the generic path: bitwise equal to forward(x)
        mean3 = (0.4, 0.5, 0.6)
        x3 = _normalize_like_reference(u8, mean3, (0.5, 0.25, 0.2))
        assert torch.equal(m.forward_u8(u8.cuda(), mean3, (0.5, 0.25, 0.2)).cpu(), m(x3.cuda()).cpu())


def test_bf16x3_in_channel_20():
    """Segment(20) (17 keypoint heat maps + image, segment.py:531-533) in the split-precision mode: the generic-cin input
    kernel writes hi + lo planes, layer1 has K = 80 per tap."""
    ref = randomized_state(20, seed=5)
    sd = ref.state_dict()
    m = _model(sd, cin=20, precision="bf16x3")
    x = torch.from_numpy(synth.batch(2, 64, 96, seed=1, channels=20))
    want = segment_forward(sd, x)
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    d = (got - want).abs()
    scale = max(1.0, float(want.abs().max()) / 3.0)
    print(f"Segment(20) bf16x3: max {float(d.max()):.3e} mean {float(d.mean()):.3e} (logit absmax {float(want.abs().max()):.2f})")
    assert float(d.max()) <= X3_MAX * scale and float(d.mean()) <= X3_MEAN * scale


def test_bf16x3_masks_agree_with_fp32_mode_masks(golden_state, golden_logits):
    """Mask-level view of the logits tolerance: thresholded masks from the tensor-core mode against masks from the fp32
    mode on the golden inputs (which the unmodified reference wrote logits for): IoU and pixel-flip rate, plus the share
    of flips that sit within the tolerance of the decision boundary (all of them must)."""
    from instancesegmentation_b200.decode import decode_masks, mask_iou
    meta = json.load(open(os.path.join(GOLDEN, "segment3_golden.json")))
    m32, m3 = _model(golden_state, precision="fp32"), _model(golden_state, precision="bf16x3")
    for name, c in meta["cases"].items():
        x = torch.from_numpy(synth.batch(c["B"], c["H"], c["W"], seed=c["seed"], kind=c["kind"])).cuda()
        with torch.no_grad():
            l32, l3 = m32(x), m3(x)
            r32, r3 = decode_masks(l32), decode_masks(l3)
        n = c["B"] * c["H"] * c["W"]
        a = torch.cat([r32.mask(k).reshape(-1) for k in range(c["B"])])
        b = torch.cat([r3.mask(k).reshape(-1) for k in range(c["B"])])
        flips = a != b
        rate = float(flips.float().mean())
        iou = mask_iou(a, b)
        scale = max(1.0, c["logit_absmax"] / 3.0)
        near = l32.reshape(-1)[flips].abs().max().item() if bool(flips.any()) else 0.0
        print(f"{name}: IoU {iou:.6f} flip rate {rate:.2e} ({int(flips.sum())} of {n}), largest |fp32 logit| at a flip {near:.2e}")
        assert iou >= 0.998 and rate <= 2e-3, (name, iou, rate)
        assert near <= X3_MAX * scale, (name, near)      # a pixel only flips when its logit is within the tolerance of 0


def test_graph_replay_sees_reloaded_weights(golden_state):
    """GraphedStep replays baked kernel arguments: a weight reload updates the packed weights in place (including the
    exact-uint8 stem variant, which is otherwise packed lazily inside isg_forward_u8), so the replay equals a fresh call."""
    from instancesegmentation_b200.pipeline import GraphedStep
    m = _model(golden_state, precision="bf16x3", max_batch=2)
    u8 = _u8_batch(2, 64, 96, seed=3).cuda()
    g = GraphedStep(m, u8, decode=False)
    g.replay()
    torch.cuda.synchronize()
    before = g.logits.clone()
    with torch.no_grad():
        assert torch.equal(before, m.forward_u8(u8))
    sd2 = {k: (v * 1.05 if v.dtype.is_floating_point and k.endswith("conv.weight") else v) for k, v in golden_state.items()}
    m.load_state_dict(sd2)
    eng = _engine(m)
    eng.load_state_dict(m.state_dict()); eng.weights_epoch = m._weights_epoch     # what the next eager call would do
    g.replay()
    torch.cuda.synchronize()
    after = g.logits.clone()
    with torch.no_grad():
        fresh = m.forward_u8(u8)
    assert float((after - before).abs().max()) > 1e-3
    assert torch.equal(after, fresh)
