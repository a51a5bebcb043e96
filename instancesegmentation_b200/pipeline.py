"""Serving pipeline around the hot path: host batches in, masks out, copies overlapped with compute.

Three CUDA streams per GPU -- host->device, compute (forward + decode), device->host -- and
``depth`` device-side slots, so that the PCIe copy of batch i+1 and the read-back of batch i-1
run while batch i is in the network.  A batch is either what the reference feeds its model
(fp32 NCHW, already normalised: train_instance.py:80-85,373) or what an image decoder produces
(uint8 NHWC; ``Normalize`` then runs inside the first kernel on the device, 4x fewer PCIe bytes).
Everything on the device runs in libisg; this file only orders streams.
"""
import numpy as np
import torch

from .decode import DecodePlan, decode_masks


class _Slot:
    def __init__(self):
        self.x = None
        self.masks = None
        self.loaded = torch.cuda.Event()
        self.consumed = torch.cuda.Event()
        self.computed = torch.cuda.Event()
        self.drained = torch.cuda.Event()
        self.res = None
        self.used = False


class Pipeline:
    """``submit(host_batch, masks_host, stats_host)`` enqueues one batch; ``finish()`` waits for all."""

    def __init__(self, model, boxes, depth=2, threshold=0.5, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5)):
        p = next(model.parameters())
        if not p.is_cuda:
            raise RuntimeError("Pipeline needs the model on a CUDA device; there is no CPU path")
        self.model, self.dev, self.threshold = model, p.device, threshold
        self.mean, self.std = tuple(mean), tuple(std)
        self.depth = depth
        self.slots = [_Slot() for _ in range(depth)]
        for sl in self.slots:
            sl.plan = DecodePlan(boxes, self.dev)       # own area / bbox outputs per slot
        self.plan = self.slots[0].plan
        self.h2d = torch.cuda.Stream(self.dev)
        self.d2h = torch.cuda.Stream(self.dev)
        self.n = 0

    def submit(self, host_batch, masks_host=None, stats_host=None):
        """``host_batch``: pinned CPU tensor, fp32 [B,C,H,W] or uint8 [B,H,W,C].  ``masks_host``
        (pinned uint8, flat) and ``stats_host`` = (area [B], bbox [B,4]) pinned int32 receive the result."""
        s = self.slots[self.n % self.depth]
        self.n += 1
        compute = torch.cuda.current_stream(self.dev)
        with torch.cuda.stream(self.h2d):
            if s.used:
                self.h2d.wait_event(s.consumed)          # the forward that read this slot's input is done
            if s.x is None or s.x.shape != host_batch.shape or s.x.dtype != host_batch.dtype:
                s.x = torch.empty(host_batch.shape, dtype=host_batch.dtype, device=self.dev)
            s.x.copy_(host_batch, non_blocking=True)
            s.loaded.record(self.h2d)
        compute.wait_event(s.loaded)
        if s.used:
            compute.wait_event(s.drained)                # the slot's masks left for the host
        if s.masks is None or s.masks.numel() < self.plan.total:
            s.masks = torch.empty(self.plan.total, dtype=torch.uint8, device=self.dev)
        with torch.no_grad():
            if s.x.dtype == torch.uint8:
                logits = self.model.forward_u8(s.x, self.mean, self.std)
            else:
                logits = self.model(s.x)
            s.consumed.record(compute)
            s.res = decode_masks(logits, threshold=self.threshold, out=s.masks, plan=s.plan)
        s.computed.record(compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(s.computed)
            if masks_host is not None:
                masks_host[:self.plan.total].copy_(s.masks[:self.plan.total], non_blocking=True)
            if stats_host is not None:
                stats_host[0].copy_(s.res.area, non_blocking=True)
                stats_host[1].copy_(s.res.bbox, non_blocking=True)
            s.drained.record(self.d2h)
        s.used = True
        return s

    def finish(self):
        """Make the current stream wait for every copy in flight (then synchronise it as usual)."""
        compute = torch.cuda.current_stream(self.dev)
        for s in self.slots:
            if s.used:
                compute.wait_event(s.drained)


def identity_boxes(batch, height, width):
    return np.tile(np.array([0, 0, width, height], np.int32), (batch, 1))
