"""Input side of infer.py on the device: per-instance crop / pad / resize of the source images to the
network size (geometry of reference train_instance.py:160-196).  Runs in libisg.so only."""
import numpy as np
import torch

from . import _lib


def instance_windows(boxes_xyxy, pad=16):
    """[N,4] instance boxes (x1,y1,x2,y2) -> crop windows (x0,y0,wk,hk): the box grown by ``pad`` pixels on
    every side (train_instance.py:165-170).  The same array is the ``boxes`` argument of ``decode_masks``."""
    b = np.asarray(boxes_xyxy, np.int64).reshape(-1, 4)
    win = np.stack([b[:, 0] - pad, b[:, 1] - pad, b[:, 2] - b[:, 0] + 2 * pad, b[:, 3] - b[:, 1] + 2 * pad], 1)
    if len(win) and (win[:, 2:] < 1).any():
        raise ValueError("empty instance box")
    return win.astype(np.int32)


def crop_resize(images, windows, inst_img, size, out=None):
    """``images``: CUDA uint8 [n_img,Hs,Ws,3]; ``windows`` int32 [N,4] (x0,y0,wk,hk) and ``inst_img`` int32 [N]
    (host arrays or CUDA tensors).  Returns uint8 [N,size[0],size[1],3], the input of ``Segment.forward_u8``."""
    if not images.is_cuda:
        raise _lib.IsgError("crop_resize runs on CUDA only; there is no CPU path")
    if images.dim() != 4 or images.shape[3] != 3 or images.dtype != torch.uint8:
        raise ValueError(f"images must be uint8 [n_img,H,W,3], got {tuple(images.shape)} {images.dtype}")
    images = images.contiguous()
    dev = images.device
    if not torch.is_tensor(windows):
        windows = torch.from_numpy(np.ascontiguousarray(np.asarray(windows, np.int32).reshape(-1, 4))).to(dev)
    if not torch.is_tensor(inst_img):
        host = np.ascontiguousarray(np.asarray(inst_img, np.int32))
        if len(host) and (host.min() < 0 or host.max() >= images.shape[0]):
            raise ValueError("inst_img out of range")
        inst_img = torch.from_numpy(host).to(dev)
    for name, t, shape_ok in (("windows", windows, windows.dim() == 2 and windows.shape[1] == 4),
                              ("inst_img", inst_img, inst_img.dim() == 1)):
        if t.dtype != torch.int32 or t.device != dev or not t.is_contiguous() or not shape_ok:
            raise TypeError(f"{name} must be a contiguous int32 tensor on {dev} ([N,4] / [N]), got {t.dtype} "
                            f"{tuple(t.shape)} on {t.device}")
    N = windows.shape[0]
    if inst_img.numel() != N:
        raise ValueError("one image index per window")
    if out is None:
        out = torch.empty((N, size[0], size[1], 3), dtype=torch.uint8, device=dev)
    if tuple(out.shape) != (N, size[0], size[1], 3) or out.dtype != torch.uint8 or not out.is_contiguous():
        raise ValueError("out must be contiguous uint8 [N,h,w,3]")
    with torch.cuda.device(dev):
        _lib.check(_lib.lib().isg_crop_resize_u8(images.data_ptr(), images.shape[0], images.shape[1], images.shape[2],
                                                 windows.data_ptr(), inst_img.data_ptr(), N, size[0], size[1],
                                                 out.data_ptr(), torch.cuda.current_stream().cuda_stream))
    return out
