"""TEST INFRASTRUCTURE ONLY -- ctypes wrapper around oracle/decode_oracle.c.

``build()`` compiles the C restatement with gcc (``-O2 -ffp-contract=off``) into
``oracle/_build/libdecode_oracle.so`` (git-ignored, travels to the GPU box).
See the header of decode_oracle.c for what it restates and its pinning status
("parity unpinned" for the decode rows: the reference has no decode).
"""
import ctypes as C
import os
import subprocess

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
_SRC = os.path.join(_HERE, "decode_oracle.c")
_OUT = os.path.join(_HERE, "_build", "libdecode_oracle.so")
_lib = None


def build(force=False):
    os.makedirs(os.path.dirname(_OUT), exist_ok=True)
    if force or not os.path.exists(_OUT) or os.path.getmtime(_OUT) < os.path.getmtime(_SRC):
        subprocess.check_call(["gcc", "-O2", "-ffp-contract=off", "-fno-fast-math", "-fPIC",
                               "-shared", "-o", _OUT, _SRC, "-lm"])
    return _OUT


def lib():
    global _lib
    if _lib is None:
        _lib = C.CDLL(build())
        _lib.oracle_sigmoid.restype = C.c_float
        _lib.oracle_sigmoid.argtypes = [C.c_float]
        _lib.oracle_prob.restype = C.c_float
        _lib.oracle_mask_iou.restype = C.c_double
    return _lib


def _p(a, t):
    return a.ctypes.data_as(C.POINTER(t))


def mask_offsets(boxes):
    """Tight packing: instance n occupies hk*wk bytes starting at off[n]; off[N] = total."""
    boxes = np.asarray(boxes, np.int32).reshape(-1, 4)
    sizes = boxes[:, 2].astype(np.int64) * boxes[:, 3].astype(np.int64)
    off = np.zeros(len(boxes) + 1, np.int64)
    np.cumsum(sizes, out=off[1:])
    return off


def sigmoid(x):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    f = lib().oracle_sigmoid
    flat_in, flat_out = x.reshape(-1), out.reshape(-1)
    for i in range(flat_in.size):
        flat_out[i] = f(float(flat_in[i]))
    return out


def decode(logits, boxes, thr=0.5, want_soft=False):
    """logits [N,h,w] fp32, boxes [N,4] (x0,y0,wk,hk) -> dict(masks, soft, area, bbox, off)."""
    logits = np.ascontiguousarray(logits, np.float32)
    boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
    N, h, w = logits.shape
    off = mask_offsets(boxes)
    masks = np.zeros(max(int(off[-1]), 1), np.uint8)
    soft = np.zeros_like(masks) if want_soft else None
    area = np.zeros(N, np.int32)
    bbox = np.zeros((N, 4), np.int32)
    lib().oracle_decode(_p(logits, C.c_float), C.c_int(N), C.c_int(h), C.c_int(w),
                        _p(boxes, C.c_int32), _p(off, C.c_int64), C.c_float(thr),
                        _p(masks, C.c_uint8), _p(soft, C.c_uint8) if want_soft else None,
                        _p(area, C.c_int32), _p(bbox, C.c_int32))
    return dict(masks=masks[:int(off[-1])], soft=None if soft is None else soft[:int(off[-1])],
                area=area, bbox=bbox, off=off)


def idmap(logits, boxes, img_first, Hs, Ws, thr=0.5):
    logits = np.ascontiguousarray(logits, np.float32)
    boxes = np.ascontiguousarray(boxes, np.int32).reshape(-1, 4)
    img_first = np.ascontiguousarray(img_first, np.int32)
    _, h, w = logits.shape
    n_img = len(img_first) - 1
    out = np.zeros((n_img, Hs, Ws), np.int32)
    lib().oracle_idmap(_p(logits, C.c_float), C.c_int(h), C.c_int(w), _p(boxes, C.c_int32),
                       _p(img_first, C.c_int32), C.c_int(n_img), C.c_int(Hs), C.c_int(Ws),
                       C.c_float(thr), _p(out, C.c_int32))
    return out


def mask_iou(a, b):
    a = np.ascontiguousarray(a, np.uint8).reshape(-1)
    b = np.ascontiguousarray(b, np.uint8).reshape(-1)
    return float(lib().oracle_mask_iou(_p(a, C.c_uint8), _p(b, C.c_uint8), C.c_int64(a.size)))


def crop_resize(src, win, inst_img, size):
    """src uint8 [n_img,Hs,Ws,C], win int32 [N,4] (x0,y0,wk,hk), inst_img int32 [N] -> uint8 [N,S_h,S_w,C]."""
    src = np.ascontiguousarray(src, np.uint8)
    win = np.ascontiguousarray(win, np.int32).reshape(-1, 4)
    inst_img = np.ascontiguousarray(inst_img, np.int32)
    _, Hs, Ws, Cc = src.shape
    out = np.zeros((len(win), size[0], size[1], Cc), np.uint8)
    lib().oracle_crop_resize(_p(src, C.c_uint8), C.c_int(Hs), C.c_int(Ws), C.c_int(Cc), _p(win, C.c_int32),
                             _p(inst_img, C.c_int32), C.c_int(len(win)), C.c_int(size[0]), C.c_int(size[1]),
                             _p(out, C.c_uint8))
    return out
