"""Keypoint heat maps of the ``Segment(20)`` variant on the device (reference train_instance.py:33-68
``keypoint2heatmaps``, consumed by ``Segment.train_batch`` at model/segment.py:531-533).  Runs in
libisg.so only."""
import torch

from . import _lib

N_PARTS = 17   # len(ORDER_PART_NAMES), train_instance.py:25-30


def keypoint_heatmaps(kp, shape, sigma=10.0, threshold=0.01, out=None):
    """``kp``: CUDA float32 [B,17,3] rows (x, y, flag), flag 1 = present and visible.
    Returns float32 [B,17,h,w] (or fills ``out``, which may be a channel slice ``t[:, 3:]`` of a
    contiguous [B,C,h,w] tensor)."""
    if not kp.is_cuda:
        raise _lib.IsgError("keypoint_heatmaps runs on CUDA only; there is no CPU path")
    if kp.dim() != 3 or kp.shape[2] != 3 or kp.dtype != torch.float32:
        raise ValueError(f"kp must be float32 [B,parts,3], got {tuple(kp.shape)} {kp.dtype}")
    kp = kp.contiguous()
    B, P, _ = kp.shape
    h, w = shape
    if out is None:
        out = torch.empty((B, P, h, w), dtype=torch.float32, device=kp.device)
    if out.shape != (B, P, h, w) or out.dtype != torch.float32 or out.device != kp.device:
        raise ValueError("out must be float32 [B,parts,h,w] on the keypoints' device")
    if B and (out.stride(3) != 1 or out.stride(2) != w or out.stride(1) != h * w):
        raise ValueError("out planes must be contiguous (a channel slice of a contiguous NCHW tensor is fine)")
    with torch.cuda.device(kp.device):
        _lib.check(_lib.lib().isg_keypoint_heatmaps(kp.data_ptr(), B, P, h, w, float(sigma), float(threshold),
                                                    out.data_ptr(), out.stride(0) if B else P * h * w,
                                                    torch.cuda.current_stream().cuda_stream))
    return out


def stack_input(image, kp, sigma=10.0, threshold=0.01):
    """The [B,3+17,H,W] input of ``Segment(20)``: ``image`` [B,3,H,W] in channels 0..2, the heat maps
    written in place into channels 3.. (the ``torch.cat`` of segment.py:531-533 without the copy of
    the maps)."""
    B, C, H, W = image.shape
    x = torch.empty((B, C + kp.shape[1], H, W), dtype=torch.float32, device=image.device)
    x[:, :C].copy_(image)
    keypoint_heatmaps(kp, (H, W), sigma, threshold, out=x[:, C:])
    return x
