"""Out-of-bounds writes of the kernels that fill caller-owned buffers, checked with guard bands: every output is a
slice of a larger buffer whose surroundings hold a sentinel that must survive (compute-sanitizer is closed on the
GPU pool this repo is developed on; scripts/sanitize.sh is the recipe for a box where it is available)."""
import numpy as np
import pytest
import torch

from instancesegmentation_b200 import Segment, synth
from instancesegmentation_b200.decode import DecodePlan, decode_masks
from instancesegmentation_b200.preprocess import crop_resize

pytestmark = pytest.mark.gpu
G = 4096   # guard elements on each side


def guarded(n, dtype, fill):
    buf = torch.full((n + 2 * G,), fill, dtype=dtype, device="cuda")
    return buf, buf[G:G + n]


def intact(buf, n, fill):
    return bool((buf[:G] == fill).all()) and bool((buf[G + n:] == fill).all())


@pytest.mark.parametrize("precision", ["fp32", "bf16", "bf16x3"])
@pytest.mark.parametrize("B,H,W", [(3, 48, 80), (1, 512, 512)])
def test_forward_writes_only_its_logits(golden_state, precision, B, H, W):
    m = Segment(3)
    m.load_state_dict(golden_state)
    m.precision, m.max_batch = precision, 2            # B = 3 also exercises a short last group
    m = m.cuda().eval()
    x = torch.from_numpy(synth.batch(B, H, W, seed=8)).cuda()
    n = B * H * W
    buf, out = guarded(n, torch.float32, -777.0)
    eng = m._engine_for(x.device, B, H, W)
    eng.forward(x, out=out.view(B, 1, H, W))
    torch.cuda.synchronize()
    assert intact(buf, n, -777.0)
    assert bool(torch.isfinite(out).all()) and not bool((out == -777.0).any())
    u8 = torch.from_numpy(np.random.default_rng(2).integers(0, 256, (B, H, W, 3), dtype=np.uint8)).cuda()
    buf, out = guarded(n, torch.float32, -777.0)
    eng.forward_u8(u8, (0.5,) * 3, (0.5,) * 3, out=out.view(B, 1, H, W))
    torch.cuda.synchronize()
    assert intact(buf, n, -777.0) and not bool((out == -777.0).any())


def test_decode_and_crop_write_only_their_outputs():
    N, h, w = 6, 96, 128
    logits = torch.from_numpy(synth.logits(N, h, w, seed=5)).cuda()
    boxes, img_first = synth.boxes(2, 3, 160, 200, seed=7)
    boxes = np.minimum(boxes, 150).astype(np.int32)
    boxes[:, 2:] = np.maximum(boxes[:, 2:], 5)
    for bx, kw in ((None, {}), (boxes, {}), (boxes, dict(img_first=img_first, canvas=(160, 200)))):
        plan = DecodePlan(np.tile(np.array([0, 0, w, h], np.int32), (N, 1)) if bx is None else bx, "cuda")
        abuf, area = guarded(N, torch.int32, -5)
        bbuf, bbox = guarded(N * 4, torch.int32, -5)
        plan.area, plan.bbox = area, bbox.view(N, 4)
        mbuf, masks = guarded(plan.total, torch.uint8, 77)
        r = decode_masks(logits, out=masks, plan=plan, **kw)
        torch.cuda.synchronize()
        assert intact(mbuf, plan.total, 77) and intact(abuf, N, -5) and intact(bbuf, N * 4, -5)
        assert int(r.area.min()) >= 0
    src = torch.from_numpy(np.random.default_rng(1).integers(0, 256, (2, 160, 200, 3), dtype=np.uint8)).cuda()
    win = torch.from_numpy(boxes).cuda()
    inst = torch.from_numpy(np.repeat(np.arange(2, dtype=np.int32), 3)).cuda()
    n = 6 * 64 * 96 * 3
    cbuf, c = guarded(n, torch.uint8, 201)
    crop_resize(src, win, inst, (64, 96), out=c.view(6, 64, 96, 3))
    torch.cuda.synchronize()
    assert intact(cbuf, n, 201)
