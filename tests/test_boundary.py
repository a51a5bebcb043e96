"""CPU tests of the drop-in boundary: state_dict compatibility, seeded init, the C-ABI
library's exported symbols and network description, and loud failure without CUDA."""
import copy
import ctypes
import hashlib
import os
import re

import pytest
import torch

from conftest import ROOT
from instancesegmentation_b200 import Segment, _lib
from test_oracle import KEYS_SHA


def test_import_path_of_the_reference():
    from model.segment import Segment as S  # train_instance.py:294
    assert S is Segment


def test_state_dict_keys_shapes_dtypes():
    m = Segment(3)
    sd = m.state_dict()
    assert len(sd) == 588
    assert hashlib.sha256("\n".join(sd.keys()).encode()).hexdigest() == KEYS_SHA
    assert sum(p.numel() for p in m.parameters()) == 257145
    assert sd["init_conv.layer1.bn.num_batches_tracked"].dtype == torch.int64
    assert tuple(sd["bottle6_1.weight"].shape) == (16, 4, 8, 8)
    assert tuple(sd["bottle4_1up.convs.1.weight"].shape) == (16, 16, 4, 4)
    assert tuple(sd["bottle5_1up.uppool.1.weight"].shape) == (16, 35, 1, 1)
    for k in ("bottle4_2.prelu.weight", "bottle4_3.prelu.weight", "bottle5_2.prelu.weight"):
        assert k in sd  # unused-but-present parameters
    assert m.init_Dim == 19 and m.bottle1_Dim == 48 and m.bottle2_Dim == 128
    assert m.export is False and m.output_mid_features is False
    assert sum(p.numel() for p in Segment(20).parameters()) == 266121


def test_golden_checkpoint_loads_strict(golden_state):
    m = Segment(3)
    m.load_state_dict(golden_state, strict=True)
    ckpt = {"branch_name": "x", "best": 0.8, "epoch": 1, "state_dict": m.state_dict(),
            "optimizer": torch.optim.Adam(m.parameters()).state_dict()}  # train_instance.py:497-503
    m2 = Segment(3)
    m2.load_state_dict(copy.deepcopy(ckpt)["state_dict"], strict=True)
    assert all(torch.equal(a, b) for a, b in zip(m.state_dict().values(), m2.state_dict().values()))


def test_seeded_init_equals_reference(reference_module):
    for cin in (3, 20):
        torch.manual_seed(11)
        want = reference_module.Segment(cin).state_dict()
        torch.manual_seed(11)
        got = Segment(cin).state_dict()
        assert list(want) == list(got)
        assert all(torch.equal(want[k], got[k]) and want[k].dtype == got[k].dtype for k in want)


def test_train_mode_equals_reference(reference_module, golden_state):
    a, b = reference_module.Segment(3), Segment(3)
    a.load_state_dict(golden_state); b.load_state_dict(golden_state)
    a.train(); b.train()
    x = torch.randn(2, 3, 64, 48)
    ya, yb = a.train_batch(x, x[:, :0]), b.train_batch(x)
    assert torch.equal(ya, yb)
    ya.mean().backward(); yb.mean().backward()
    ga = dict(a.named_parameters()); gb = dict(b.named_parameters())
    assert all((ga[k].grad is None) == (gb[k].grad is None) for k in ga)
    assert all(torch.equal(ga[k].grad, gb[k].grad) for k in ga if ga[k].grad is not None)
    assert torch.equal(a.state_dict()["bottle3_1.convs.0.bn.running_var"],
                       b.state_dict()["bottle3_1.convs.0.bn.running_var"])


def test_size_rule_and_no_cpu_fallback():
    m = Segment(3).eval()
    with pytest.raises(ValueError):
        m(torch.zeros(1, 3, 500, 512))
    with pytest.raises(ValueError):
        m(torch.zeros(1, 4, 64, 64))
    with pytest.raises(RuntimeError, match="no CPU fallback"):
        with torch.no_grad():
            m(torch.zeros(1, 3, 64, 64))


def test_library_exports_every_declared_symbol():
    header = open(os.path.join(ROOT, "include", "isg.h")).read()
    declared = set(re.findall(r"ISG_API[^;(]*?\b(isg_\w+)\s*\(", header))
    assert declared == set(_lib.SIGNATURES), declared ^ set(_lib.SIGNATURES)
    raw = ctypes.CDLL(_lib.LIB_PATH)
    for name in declared:
        assert hasattr(raw, name), name
    assert _lib.lib().isg_version() >= 100


def test_library_network_description_matches_module():
    l = _lib.lib()
    for cin in (3, 20):
        sd = Segment(cin).state_dict()
        assert l.isg_state_count(cin) == len(sd)
        for i, (k, v) in enumerate(sd.items()):
            assert l.isg_state_name(cin, i).decode() == k
            assert l.isg_state_numel(cin, i) == v.numel()
            assert l.isg_state_is_int64(cin, i) == int(v.dtype == torch.int64)
    assert l.isg_state_name(3, 588) is None


def test_library_rejects_bad_arguments_without_gpu():
    l = _lib.lib()
    h = ctypes.c_void_p()
    assert l.isg_plan_create(0, 3, 500, 512, 1, 0, ctypes.byref(h)) == -1
    assert b"multiples of 16" in l.isg_last_error()
    assert l.isg_plan_create(0, 3, 64, 64, 1, 7, ctypes.byref(h)) == -1
    assert l.isg_forward(None, None, 1, None, None) == -1
    # entry points added for the ragged decode and the input side: argument checks come before any CUDA call
    assert l.isg_decode_tile_rows() == 16
    assert l.isg_decode_tiled(None, 0, 1, 64, 64, None, None, None, 1, 0.5, None, None, None, None, None, 0, 0, 0, None, None) == -1
    assert b"null argument" in l.isg_last_error()
    assert l.isg_decode_tiled(None, 0, 0, 64, 64, None, None, None, 0, 0.5, None, None, None, None, None, 0, 0, 0, None, None) == 0
    buf = (ctypes.c_uint8 * 64)()
    pbuf = ctypes.cast(buf, ctypes.c_void_p)
    assert l.isg_crop_resize_u8(pbuf, 1, 8, 8, pbuf, pbuf, 1, 16, 24, pbuf, None) == -1
    assert b"multiple of 16" in l.isg_last_error()
    assert l.isg_keypoint_heatmaps(pbuf, 1, 17, 16, 18, 10.0, 0.01, pbuf, 17 * 16 * 18, None) == -1
    assert b"multiple of 4" in l.isg_last_error()
    assert l.isg_keypoint_heatmaps(pbuf, 1, 17, 16, 16, 10.0, 1.5, pbuf, 17 * 16 * 16, None) == -1
    assert b"threshold" in l.isg_last_error()
    # training-step entry points (SURVEY 8f-1 slice)
    assert l.isg_bce_workspace_bytes() >= 8 * 1024 and l.isg_bn_max_split() >= 1
    assert l.isg_bce_sigmoid(pbuf, 0, pbuf, 0, None, pbuf, pbuf, None) == -1
    assert b"empty batch" in l.isg_last_error()
    assert l.isg_bce_sigmoid(pbuf, 0, None, 4, None, pbuf, pbuf, None) == -1
    bn_fwd = lambda batch, mom, act, slope, ns: l.isg_bn_train_forward(
        pbuf, 0, batch, 4, 4, None, None, None, None, None, mom, 1e-5, act, slope, ns, pbuf, pbuf, pbuf, pbuf, None)
    assert bn_fwd(0, 0.1, 0, None, 0) == -1
    assert b"bad sizes" in l.isg_last_error()
    assert bn_fwd(1, 1.5, 0, None, 0) == -1
    assert b"momentum" in l.isg_last_error()
    assert bn_fwd(1, 0.1, 3, None, 0) == -1
    assert b"act must be" in l.isg_last_error()
    assert bn_fwd(1, 0.1, 2, pbuf, 3) == -1
    assert b"slopes" in l.isg_last_error()
    assert l.isg_bn_train_backward(pbuf, None, 0, 1, 4, 4, None, None, pbuf, pbuf, 0, None, 0, pbuf, None, None, None,
                                   pbuf, None) == -1
    assert l.isg_dwconv_forward(pbuf, 0, 1, 4, 4, 4, pbuf, None, 3, 3, 0, 1, 1, 1, 0, pbuf, None) == -1
    assert b"2*pad == dil*(k-1)" in l.isg_last_error()
    assert l.isg_dwconv_wgrad(pbuf, None, 0, 1, 4, 4, 4, 3, 3, 1, 1, 1, 1, pbuf, None, None, None) == -1
    assert l.isg_dwconv_wgrad_workspace_bytes(16) == 8 * 16 * 26 * 4
    assert l.isg_convt_k2s_forward(pbuf, 0, 1, 16, 4, 8, 8, 8, 3, 2, pbuf, None, pbuf, None) == -1
    assert b"kernel == 2*stride" in l.isg_last_error()


def test_infer_call_surface():
    import infer  # reference infer.py:12-29
    assert infer.path_decompose("/a/b/img.01.png") == ("/a/b", "img.01", "png")
    args = infer.parse_args(["-i", "in", "-o", "out", "--continue-test"])
    assert (args.test_image_dir, args.output_dir, args.continue_test) == ("in", "out", True)
    with pytest.raises(SystemExit):
        infer.parse_args(["-i", "in"])
