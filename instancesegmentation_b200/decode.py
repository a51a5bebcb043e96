"""Mask decode -- the body the reference's infer.py never got (infer.py:32-36): logits ->
sigmoid (model/segment.py:534) -> bilinear resize to each instance's source box ->
threshold -> per-instance binary masks, areas, boxes, optional soft masks
(train_instance.py:398-399) and the per-image id map.  Runs in libisg.so only."""
import ctypes as C
from dataclasses import dataclass
from typing import Optional

import numpy as np
import torch

from . import _lib


@dataclass
class DecodeResult:
    masks: torch.Tensor            # flat uint8 (0/1) on the device
    soft: Optional[torch.Tensor]   # flat uint8 trunc(p*255), same layout, or None
    offsets: np.ndarray            # int64 [N+1] host byte offsets (instance n starts at offsets[n])
    boxes: np.ndarray              # int32 [N,4] host (x0,y0,wk,hk)
    area: torch.Tensor             # int32 [N] device
    bbox: torch.Tensor             # int32 [N,4] device (xmin,ymin,xmax+1,ymax+1), zeros if empty
    idmap: Optional[torch.Tensor] = None  # int32 [n_img,Hs,Ws] device

    def mask(self, n):
        wk, hk = int(self.boxes[n, 2]), int(self.boxes[n, 3])
        o = int(self.offsets[n])
        return self.masks[o:o + hk * wk].view(hk, wk)

    def soft_mask(self, n):
        wk, hk = int(self.boxes[n, 2]), int(self.boxes[n, 3])
        o = int(self.offsets[n])
        return self.soft[o:o + hk * wk].view(hk, wk)


def mask_offsets(boxes, align=16):
    """Byte offset of every instance's hk*wk mask; starts are `align`-byte aligned so the
    kernel's 4-byte stores are always used (align=1 gives the oracle's tight packing)."""
    boxes = np.asarray(boxes, np.int32).reshape(-1, 4)
    sizes = boxes[:, 2].astype(np.int64) * boxes[:, 3].astype(np.int64)
    off = np.zeros(len(boxes) + 1, np.int64)
    for i, s in enumerate(sizes):
        off[i + 1] = (off[i] + s + align - 1) // align * align
    return off


class DecodePlan:
    """Device-resident geometry of a decode that is repeated with the same boxes (serving loop,
    bench): box table, mask offsets and the area/bbox outputs are uploaded / allocated once, so a
    call is exactly one kernel launch with no host<->device traffic."""

    def __init__(self, boxes, device, align=16):
        self.boxes = np.ascontiguousarray(np.asarray(boxes, np.int32).reshape(-1, 4))
        if len(self.boxes) and (self.boxes[:, 2:] < 1).any():
            raise ValueError("box width/height must be >= 1")
        self.offsets = mask_offsets(self.boxes, align)
        self.device = torch.device(device)
        n = len(self.boxes)
        self.boxes_d = torch.from_numpy(self.boxes).to(self.device)
        self.off_d = torch.from_numpy(self.offsets[:-1].copy()).to(self.device)
        self.area = torch.empty(max(n, 1), dtype=torch.int32, device=self.device)[:n]
        self.bbox = torch.empty((max(n, 1), 4), dtype=torch.int32, device=self.device)[:n]
        self.max_w = int(self.boxes[:, 2].max()) if n else 1
        self.max_h = int(self.boxes[:, 3].max()) if n else 1
        self.total = max(int(self.offsets[-1]), 1)
        # tiled kernel (isg_decode_tiled): exclusive prefix of the instances' tile counts
        rows = _lib.lib().isg_decode_tile_rows()
        hk = np.where(self.boxes[:, 2] > 0, self.boxes[:, 3], 0).astype(np.int64) if n else np.zeros(0, np.int64)
        tf = np.zeros(n + 1, np.int64)
        np.cumsum((np.maximum(hk, 0) + rows - 1) // rows, out=tf[1:])
        if tf[-1] >= 2 ** 31:
            raise ValueError("too many mask pixels for one decode call")
        self.n_tiles = int(tf[-1])
        self.tile_first_d = torch.from_numpy(tf.astype(np.int32)).to(self.device)
        # this plan's own accumulator workspace (isg_decode_bind_workspace): decodes of different plans may overlap on
        # the GPU (pipeline slots, several streams); zeroed once, every launch leaves it zeroed
        self.ws_inst = max(n, 1)
        self.ws = torch.zeros((_lib.lib().isg_decode_workspace_bytes(self.ws_inst) + 15) // 16 * 4, dtype=torch.int32,
                              device=self.device)
        self._idmap_tables, self._identity = {}, {}

    def is_identity(self, h, w):
        if (h, w) not in self._identity:
            b = self.boxes
            self._identity[(h, w)] = bool(len(b)) and bool((b[:, 2] == w).all() and (b[:, 3] == h).all())
        return self._identity[(h, w)]

    def idmap_tables(self, img_first):
        """(img_first on the device, image of every instance on the device, max instances per image)"""
        first = np.ascontiguousarray(np.asarray(img_first, np.int32))
        key = first.tobytes()
        if key not in self._idmap_tables:
            if first[0] != 0 or first[-1] != len(self.boxes) or (np.diff(first) < 0).any():
                raise ValueError("img_first must rise from 0 to the number of instances")
            counts = np.diff(first)
            inst_img = np.repeat(np.arange(len(counts), dtype=np.int32), counts)
            self._idmap_tables[key] = (torch.from_numpy(first).to(self.device),
                                       torch.from_numpy(inst_img).to(self.device),
                                       int(counts.max()) if len(counts) else 0)
        return self._idmap_tables[key]


def decode_masks(logits, boxes=None, threshold=0.5, want_soft=False, img_first=None, canvas=None,
                 align=16, out=None, plan=None):
    """``logits``: CUDA fp32/bf16 [N,h,w] or [N,1,h,w].  ``boxes``: int32 [N,4] (x0,y0,wk,hk) on
    the host (None = identity: every map decoded at its own size at the canvas origin).
    ``img_first`` (int32 [n_img+1]) + ``canvas`` (Hs, Ws) additionally produce the id map."""
    if not logits.is_cuda:
        raise _lib.IsgError("decode_masks runs on CUDA only; there is no CPU path")
    if logits.dim() == 4:
        if logits.shape[1] != 1:
            raise ValueError("logits must have one channel")
        logits = logits[:, 0]
    if logits.dtype not in (torch.float32, torch.bfloat16):
        raise TypeError(f"logits must be float32 or bfloat16, got {logits.dtype}")
    logits = logits.contiguous()
    N, h, w = logits.shape
    dev = logits.device
    if plan is None:
        if boxes is None:
            boxes = np.tile(np.array([0, 0, w, h], np.int32), (N, 1))
        plan = DecodePlan(boxes, dev, align)
    boxes, off, total = plan.boxes, plan.offsets, plan.total
    if len(boxes) != N:
        raise ValueError(f"{N} logit maps but {len(boxes)} boxes")
    l = _lib.lib()
    with torch.cuda.device(dev):
        st = torch.cuda.current_stream().cuda_stream
        if out is not None:
            masks = out
            if masks.numel() < total or masks.dtype != torch.uint8:
                raise ValueError("out buffer too small or not uint8")
        else:
            masks = torch.empty(total, dtype=torch.uint8, device=dev)
        soft = torch.empty(total, dtype=torch.uint8, device=dev) if want_soft else None
        area, bbox, boxes_d, off_d = plan.area, plan.bbox, plan.boxes_d, plan.off_d
        bf = int(logits.dtype == torch.bfloat16)
        idmap = first_d = inst_img_d = None
        per_img = 0
        if img_first is not None:
            if canvas is None:
                raise ValueError("canvas=(Hs, Ws) is required with img_first")
            first_d, inst_img_d, per_img = plan.idmap_tables(img_first)
            n_img = first_d.numel() - 1
            idmap = torch.empty((n_img, canvas[0], canvas[1]), dtype=torch.int32, device=dev)
        identity = plan.is_identity(h, w)
        _lib.check(l.isg_decode_bind_workspace(plan.ws.data_ptr(), plan.ws_inst))
        fuse_ids = bool(idmap is not None and N and per_img <= 32 and not want_soft)
        if N and not want_soft and (fuse_ids or not identity):
            # ragged sizes: balanced tile list; the id map rides along as bit claims when it fits a word
            _lib.check(l.isg_decode_tiled(logits.data_ptr(), bf, N, h, w, boxes_d.data_ptr(), off_d.data_ptr(),
                                          plan.tile_first_d.data_ptr(), plan.n_tiles, float(threshold),
                                          masks.data_ptr(), area.data_ptr(), bbox.data_ptr(),
                                          inst_img_d.data_ptr() if fuse_ids else None,
                                          first_d.data_ptr() if fuse_ids else None,
                                          n_img if fuse_ids else 0, canvas[0] if fuse_ids else 0,
                                          canvas[1] if fuse_ids else 0,
                                          idmap.data_ptr() if fuse_ids else None, st))
        elif N:
            _lib.check(l.isg_decode(logits.data_ptr(), bf, N, h, w,
                                    boxes_d.data_ptr(), off_d.data_ptr(), plan.max_w,
                                    plan.max_h, float(threshold), masks.data_ptr(),
                                    soft.data_ptr() if want_soft else None, area.data_ptr(),
                                    bbox.data_ptr(), st))
        if idmap is not None and not fuse_ids:
            _lib.check(l.isg_idmap(logits.data_ptr(), bf, h, w,
                                   boxes_d.data_ptr(), first_d.data_ptr(), n_img, canvas[0], canvas[1],
                                   float(threshold), idmap.data_ptr(), st))
        l.isg_decode_bind_workspace(None, 0)
    return DecodeResult(masks, soft, off, boxes, area, bbox, idmap)


def pack_mask_bits(masks, n=None, out=None):
    """``masks``: flat uint8 0/1 CUDA tensor (``DecodeResult.masks``) -> uint8 tensor of ``(n + 7) // 8`` bytes with one
    bit per pixel (bit i of byte j = masks[8 j + i], ``numpy.packbits(..., bitorder="little")``): 8x fewer bytes to copy
    to the host.  ``numpy.unpackbits(bits, bitorder="little")[:n]`` restores the bytes."""
    if masks.dtype != torch.uint8 or not masks.is_cuda:
        raise TypeError("pack_mask_bits needs a uint8 CUDA tensor")
    masks = masks.contiguous().view(-1)
    n = masks.numel() if n is None else int(n)
    if n > masks.numel():
        raise ValueError("n exceeds the mask buffer")
    nb = (n + 7) // 8
    if out is None:
        out = torch.empty(nb + (nb & 1), dtype=torch.uint8, device=masks.device)
    if out.dtype != torch.uint8 or out.numel() < nb:
        raise ValueError("out must be uint8 with at least (n + 7) // 8 bytes")
    with torch.cuda.device(masks.device):
        _lib.check(_lib.lib().isg_pack_mask_bits(masks.data_ptr(), n, out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out[:nb]


def mask_iou(a, b, threshold=1):
    """intersection / union of two uint8 CUDA masks (train_instance.py:402-403, where both are ``tensor2mask``
    soft bytes 0-255): a byte is set when it is >= ``threshold`` (an int, or a pair for a and b).  The default 1 =
    any nonzero byte, numpy's logical view of soft bytes; 128 = probability >= 0.5.  0 when both masks are empty."""
    if a.dtype != torch.uint8 or b.dtype != torch.uint8:
        raise TypeError(f"mask_iou needs uint8 masks, got {a.dtype} and {b.dtype}")
    if not (a.is_cuda and b.is_cuda):
        raise _lib.IsgError("mask_iou runs on CUDA only; there is no CPU path")
    a, b = a.contiguous().view(-1), b.contiguous().view(-1)
    if a.numel() != b.numel():
        raise ValueError("mask sizes differ")
    ta, tb = (threshold, threshold) if isinstance(threshold, int) else threshold
    cnt = torch.zeros(2, dtype=torch.int64, device=a.device)
    with torch.cuda.device(a.device):
        _lib.check(_lib.lib().isg_mask_iou_counts_thr(a.data_ptr(), b.data_ptr(), a.numel(), int(ta), int(tb),
                                                      cnt.data_ptr(), torch.cuda.current_stream().cuda_stream))
    inter, union = cnt.tolist()
    return inter / union if union else 0.0
