"""GPU parity tests of the mask decode: bit-exact masks / soft bytes / area / bbox / ids against
oracle/decode_oracle.c on the same inputs, plus size-independent properties at full size."""
import numpy as np
import pytest
import torch

from instancesegmentation_b200 import synth
from instancesegmentation_b200 import _lib as _lib_mod
from instancesegmentation_b200.decode import decode_masks, mask_iou
from oracle import decode_oracle as D

pytestmark = pytest.mark.gpu


def _compare(L, boxes, thr=0.5, soft=False, align=16):
    r = decode_masks(torch.from_numpy(L).cuda(), boxes, threshold=thr, want_soft=soft, align=align)
    o = D.decode(L, boxes, thr=thr, want_soft=soft)
    masks = r.masks.cpu().numpy()
    for n in range(len(boxes)):
        want = o["masks"][o["off"][n]:o["off"][n + 1]]
        got = masks[r.offsets[n]:r.offsets[n] + len(want)]
        assert np.array_equal(got, want), (n, int((got != want).sum()))
        if soft:
            ws = o["soft"][o["off"][n]:o["off"][n + 1]]
            gs = r.soft.cpu().numpy()[r.offsets[n]:r.offsets[n] + len(ws)]
            assert np.array_equal(gs, ws), (n, int((gs != ws).sum()))
    assert np.array_equal(r.area.cpu().numpy(), o["area"])
    assert np.array_equal(r.bbox.cpu().numpy(), o["bbox"])
    return r, o


def test_identity_size_bit_exact():
    L = synth.logits(3, 64, 96, seed=1)
    L[0, :2, :8] = [[0.0, 1e-9, 1e-7, 6e-8, -1e-9, 5.9e-8, 6.1e-8, 3e-8]] * 2  # around sigmoid == 0.5
    L[1, 0, :4] = [np.nan, np.inf, -np.inf, 1e-38]
    _compare(L, np.array([[0, 0, 96, 64]] * 3, np.int32), soft=True)
    _compare(L, np.array([[0, 0, 96, 64]] * 3, np.int32))


def test_ragged_boxes_up_and_down_sampling():
    L = synth.logits(6, 64, 80, seed=2)
    boxes = np.array([[0, 0, 160, 128], [5, 7, 37, 23], [1, 2, 80, 64], [9, 9, 81, 63],
                      [0, 3, 7, 200], [4, 4, 333, 5]], np.int32)
    for align in (16, 1):     # aligned (4-byte stores) and tight packing (byte stores)
        _compare(L, boxes, align=align)
    _compare(L, boxes, soft=True)
    _compare(L, boxes, thr=0.3)
    _compare(L, boxes, thr=0.9)


def test_degenerate_masks():
    L = np.stack([np.full((32, 32), -9, np.float32), np.full((32, 32), 9, np.float32),
                  np.where(np.arange(1024).reshape(32, 32) == 517, 6, -6).astype(np.float32)])
    boxes = np.array([[3, 4, 32, 32], [0, 0, 50, 20], [10, 10, 32, 32]], np.int32)
    r, _ = _compare(L, boxes)
    assert r.area.tolist() == [0, 1000, 1]
    assert r.bbox.tolist() == [[0, 0, 0, 0], [0, 0, 50, 20], [10 + 5, 10 + 16, 10 + 6, 10 + 17]]
    r1, _ = _compare(L[:1], np.array([[0, 0, 1, 1]], np.int32))
    assert r1.masks.numel() >= 1
    empty = decode_masks(torch.zeros((0, 16, 16), device="cuda"))
    assert empty.area.numel() == 0


def test_many_overlapping_instances_and_idmap():
    n_img, per = 2, 24
    boxes, first = synth.boxes(n_img, per, 96, 128, seed=7, lo=12, hi=90)
    L = synth.logits(n_img * per, 32, 48, seed=3)
    r = decode_masks(torch.from_numpy(L).cuda(), boxes, img_first=first, canvas=(96, 128))
    o = D.decode(L, boxes)
    assert np.array_equal(r.area.cpu().numpy(), o["area"])
    want = D.idmap(L, boxes, first, 96, 128)
    got = r.idmap.cpu().numpy()
    assert np.array_equal(got, want), int((got != want).sum())
    assert got.max() > per  # ids are 1-based global instance indices


def test_idmap_more_than_32_instances_per_image_uses_gather_kernel():
    n_img, per = 1, 40
    boxes, first = synth.boxes(n_img, per, 64, 80, seed=11, lo=8, hi=60)
    L = synth.logits(n_img * per, 32, 32, seed=12)
    r = decode_masks(torch.from_numpy(L).cuda(), boxes, img_first=first, canvas=(64, 80))
    assert np.array_equal(r.idmap.cpu().numpy(), D.idmap(L, boxes, first, 64, 80))
    o = D.decode(L, boxes)
    assert np.array_equal(r.area.cpu().numpy(), o["area"]) and np.array_equal(r.bbox.cpu().numpy(), o["bbox"])


def test_tiled_decode_boxes_off_canvas_nan_and_thresholds():
    """ragged sizes through isg_decode_tiled: boxes hanging over the canvas edge, NaN/inf taps, thresholds
    next to 0/1 (the logit shortcut is off there), tight packing (byte stores)"""
    n_img, per = 3, 7
    boxes, first = synth.boxes(n_img, per, 72, 88, seed=21, lo=5, hi=70)
    boxes[0, :2] = [-9, -4]            # hangs over the top-left corner
    boxes[5, :2] = [70, 60]            # and over the bottom-right one
    L = synth.logits(n_img * per, 40, 56, seed=22)
    L[2, 3:6, 4:9] = np.nan
    L[3, 10, :] = np.inf
    L[4, :, 7] = -np.inf
    L[6, 0:8, 0:8] = [[0.0, 1e-9, 1e-7, 6e-8, -1e-9, 5.9e-8, 6.1e-8, 3e-8]] * 8
    for thr in (0.5, 0.3, 0.9995, 2e-4):
        for align in (16, 1):
            _compare(L, boxes, thr=thr, align=align)
        r = decode_masks(torch.from_numpy(L).cuda(), boxes, threshold=thr, img_first=first, canvas=(72, 88))
        want = D.idmap(L, boxes, first, 72, 88, thr=thr)
        got = r.idmap.cpu().numpy()
        assert np.array_equal(got, want), (thr, int((got != want).sum()))


def test_tiled_decode_full_size_balanced_and_deterministic():
    """cfg4-like mix of box sizes at 1024x1024: statistics agree with the masks, two runs agree, and the
    id map only names instances whose mask is set at that pixel"""
    n_img, per, S = 2, 20, 1024
    boxes, first = synth.boxes(n_img, per, S, S, seed=7, lo=96, hi=640)
    L = torch.from_numpy(synth.logits(n_img * per, S, S, seed=5)).cuda()
    r = decode_masks(L, boxes, img_first=first, canvas=(S, S))
    r2 = decode_masks(L, boxes, img_first=first, canvas=(S, S))
    assert torch.equal(r.idmap, r2.idmap) and torch.equal(r.bbox, r2.bbox) and torch.equal(r.area, r2.area)
    assert all(torch.equal(r.mask(n), r2.mask(n)) for n in range(n_img * per))   # (alignment gaps are never written)
    ids = r.idmap.cpu().numpy()
    cover = np.zeros_like(ids)
    for n in range(n_img * per):
        m = r.mask(n)
        x0, y0, wk, hk = (int(v) for v in boxes[n])
        assert int(r.area[n]) == int(m.sum())
        if int(r.area[n]):
            ys, xs = torch.nonzero(m, as_tuple=True)
            assert r.bbox[n].tolist() == [x0 + int(xs.min()), y0 + int(ys.min()), x0 + int(xs.max()) + 1, y0 + int(ys.max()) + 1]
        img = int(np.searchsorted(first, n, side="right") - 1)
        mm = m.cpu().numpy()
        ya, yb, xa, xb = max(y0, 0), min(y0 + hk, S), max(x0, 0), min(x0 + wk, S)
        sub = mm[ya - y0:yb - y0, xa - x0:xb - x0]
        cover[img, ya:yb, xa:xb] |= sub
        named = ids[img, ya:yb, xa:xb] == n + 1
        assert not (named & (sub == 0)).any()
    assert np.array_equal(ids != 0, cover != 0)
    # a 96-row slice of the first instance against the scalar oracle
    o = D.decode(L[:1].cpu().numpy(), boxes[:1])
    assert np.array_equal(r.mask(0).cpu().numpy().ravel(), o["masks"][:o["off"][1]])


def test_bf16_logits():
    L = torch.from_numpy(synth.logits(2, 48, 64, seed=9)).to(torch.bfloat16)
    boxes = np.array([[0, 0, 64, 48], [2, 2, 100, 70]], np.int32)
    r = decode_masks(L.cuda(), boxes, want_soft=True)
    o = D.decode(L.float().numpy(), boxes, want_soft=True)
    for n in range(2):
        want = o["masks"][o["off"][n]:o["off"][n + 1]]
        assert np.array_equal(r.masks.cpu().numpy()[r.offsets[n]:r.offsets[n] + len(want)], want)
    assert np.array_equal(r.area.cpu().numpy(), o["area"])


def test_full_size_properties():
    """BASELINE config sizes, checked through properties (the scalar oracle is too slow here)."""
    N, h, w = 16, 512, 512
    L = torch.from_numpy(synth.logits(N, h, w, seed=4)).cuda()
    r = decode_masks(L)
    m = r.masks.view(N, h, w)
    assert torch.equal(m.bool(), torch.sigmoid(L) > 0.5) or (m.bool() != (torch.sigmoid(L) > 0.5)).sum() <= 2
    assert torch.equal(r.area, m.sum(dim=(1, 2), dtype=torch.int32))
    again = decode_masks(L)                                   # idempotent / deterministic
    assert torch.equal(again.masks, r.masks) and torch.equal(again.bbox, r.bbox)
    neg = decode_masks(-L)                                    # complement away from ties
    assert int((neg.masks.view(N, h, w) + m != 1).sum()) <= int((L == 0).sum()) + 2
    for n in range(N):                                        # bbox is tight
        ys, xs = torch.nonzero(m[n], as_tuple=True)
        assert r.bbox[n].tolist() == [int(xs.min()), int(ys.min()), int(xs.max()) + 1, int(ys.max()) + 1]
    up = decode_masks(L[:2], np.array([[0, 0, 1024, 1024]] * 2, np.int32))   # 2x upsampling
    o = D.decode(L[:1, :64, :64].cpu().contiguous().numpy(), np.array([[0, 0, 128, 128]], np.int32))
    sub = decode_masks(L[:1, :64, :64].contiguous(), np.array([[0, 0, 128, 128]], np.int32))
    assert np.array_equal(sub.masks.cpu().numpy()[:128 * 128], o["masks"])
    assert abs(int(up.area[0]) - 4 * int(r.area[0])) <= 0.02 * 4 * int(r.area[0]) + 64
    assert mask_iou(r.masks, again.masks) == 1.0


def test_mask_iou_counts_against_oracle_on_random_and_soft_masks():
    """isg_mask_iou_counts(_thr) vs oracle_mask_iou (nonzero = set): 0/1 masks, the reference's 0-255 soft bytes
    (train_instance.py:398-403 feeds tensor2mask outputs), thresholded soft bytes, odd sizes, empty masks."""
    from oracle import decode_oracle as D
    rng = np.random.default_rng(3)
    for n in (1, 31, 4097, 480 * 480, 3 * 512 * 512 + 5):
        a = (rng.random(n) < 0.4).astype(np.uint8)
        b = (rng.random(n) < 0.55).astype(np.uint8)
        assert mask_iou(torch.from_numpy(a).cuda(), torch.from_numpy(b).cuda()) == D.mask_iou(a, b)
        sa = rng.integers(0, 256, n, dtype=np.uint8) * (rng.random(n) < 0.7).astype(np.uint8)   # soft bytes with exact zeros
        sb = rng.integers(0, 256, n, dtype=np.uint8) * (rng.random(n) < 0.6).astype(np.uint8)
        ta, tb = torch.from_numpy(sa).cuda(), torch.from_numpy(sb).cuda()
        assert mask_iou(ta, tb) == D.mask_iou(sa, sb)                                            # nonzero rule
        assert mask_iou(ta, tb, 128) == D.mask_iou((sa >= 128).astype(np.uint8), (sb >= 128).astype(np.uint8))
        assert mask_iou(ta, tb, (200, 3)) == D.mask_iou((sa >= 200).astype(np.uint8), (sb >= 3).astype(np.uint8))
    z = torch.zeros(1000, dtype=torch.uint8, device="cuda")
    assert mask_iou(z, z) == 0.0
    with pytest.raises(TypeError):
        mask_iou(z.float(), z)
    with pytest.raises(_lib_mod.IsgError):
        mask_iou(z, z, 0)


def test_decode_workspaces_allow_concurrent_streams_and_default_path_still_works():
    """Every DecodePlan owns its accumulator workspace (isg_decode_bind_workspace): decodes on two streams may overlap.
    The unbound default (the library's per-device arrays) is reached through the raw C ABI and gives the same result."""
    import ctypes as C
    from instancesegmentation_b200.decode import DecodePlan
    N, h, w = 40, 96, 128
    la = torch.from_numpy(synth.logits(N, h, w, seed=11)).cuda()
    lb = torch.from_numpy(synth.logits(N, h, w, seed=12)).cuda()
    boxes, _ = synth.boxes(2, 20, 300, 300, seed=3)
    ref_a, ref_b = decode_masks(la, boxes), decode_masks(lb, boxes)
    ident_a = decode_masks(la)
    torch.cuda.synchronize()
    pa, pb = DecodePlan(boxes, "cuda"), DecodePlan(boxes, "cuda")
    pi = DecodePlan(np.tile(np.array([0, 0, w, h], np.int32), (N, 1)), "cuda")
    assert pa.ws.data_ptr() != pb.ws.data_ptr()
    sa, sb, sc = torch.cuda.Stream(), torch.cuda.Stream(), torch.cuda.Stream()
    for _ in range(20):
        with torch.cuda.stream(sa):
            ra = decode_masks(la, plan=pa)
        with torch.cuda.stream(sb):
            rb = decode_masks(lb, plan=pb)
        with torch.cuda.stream(sc):
            ri = decode_masks(la, plan=pi)
    torch.cuda.synchronize()
    def same(r, ref):   # per instance: the flat buffer has alignment padding between masks that nobody writes
        return (all(torch.equal(r.mask(n), ref.mask(n)) for n in range(N)) and torch.equal(r.area, ref.area)
                and torch.equal(r.bbox, ref.bbox))
    assert same(ra, ref_a) and same(rb, ref_b) and same(ri, ident_a)
    assert int(pa.ws.abs().sum()) == 0 and int(pi.ws.abs().sum()) == 0        # left zeroed for the next launch
    # default path: no workspace bound
    l = _lib_mod.lib()
    masks = torch.zeros(pi.total, dtype=torch.uint8, device="cuda")
    area = torch.zeros(N, dtype=torch.int32, device="cuda")
    bbox = torch.zeros((N, 4), dtype=torch.int32, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib_mod.check(l.isg_decode_bind_workspace(None, 0))
    _lib_mod.check(l.isg_decode(la.data_ptr(), 0, N, h, w, pi.boxes_d.data_ptr(), pi.off_d.data_ptr(), w, h, 0.5,
                                masks.data_ptr(), None, area.data_ptr(), bbox.data_ptr(), st))
    torch.cuda.synchronize()
    assert all(torch.equal(masks[int(pi.offsets[n]):int(pi.offsets[n]) + h * w].view(h, w), ident_a.mask(n)) for n in range(N))
    assert torch.equal(area, ident_a.area) and torch.equal(bbox, ident_a.bbox)
    small = torch.zeros(16, dtype=torch.int32, device="cuda")
    _lib_mod.check(l.isg_decode_bind_workspace(small.data_ptr(), 1))
    assert l.isg_decode(la.data_ptr(), 0, N, h, w, pi.boxes_d.data_ptr(), pi.off_d.data_ptr(), w, h, 0.5,
                        masks.data_ptr(), None, area.data_ptr(), bbox.data_ptr(), st) == -1     # 40 instances > workspace of 1
    _lib_mod.check(l.isg_decode_bind_workspace(None, 0))


def test_packed_mask_bits_equal_the_byte_masks():
    """isg_pack_mask_bits: 1 bit per pixel, bit-exact against the byte masks for every length (tails of 1..15 bytes),
    and through Pipeline(packed=True)."""
    from instancesegmentation_b200.decode import pack_mask_bits
    rng = np.random.default_rng(5)
    for n in (1, 7, 8, 9, 15, 16, 17, 31, 4096, 4099, 3 * 512 * 512 + 13):
        m = (rng.random(n) < 0.5).astype(np.uint8)
        buf = torch.zeros((n + 15) // 16 * 16, dtype=torch.uint8, device="cuda")
        buf[:n] = torch.from_numpy(m).cuda()
        bits = pack_mask_bits(buf, n).cpu().numpy()
        assert np.array_equal(bits, np.packbits(m, bitorder="little")), n
        assert np.array_equal(np.unpackbits(bits, bitorder="little")[:n], m)
    logits = torch.from_numpy(synth.logits(5, 96, 128, seed=9)).cuda()
    r = decode_masks(logits)
    bits = pack_mask_bits(r.masks, r.offsets[-1]).cpu().numpy()
    back = np.unpackbits(bits, bitorder="little")
    for k in range(5):
        o = int(r.offsets[k])
        assert np.array_equal(back[o:o + 96 * 128].reshape(96, 128), r.mask(k).cpu().numpy())
