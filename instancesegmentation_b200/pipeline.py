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

from .decode import DecodePlan, decode_masks, pack_mask_bits


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
        self.graph = None


class Pipeline:
    """``submit(host_batch, masks_host, stats_host)`` enqueues one batch; ``finish()`` waits for all."""

    def __init__(self, model, boxes, depth=2, threshold=0.5, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5), graphs=False,
                 packed=False):
        """``graphs=True``: forward + decode of every slot is captured once (GraphedStep) and replayed.
        ``packed=True``: the masks leave the device as 1 bit per pixel (``pack_mask_bits``; ``masks_host`` then receives
        ``(plan.total + 7) // 8`` bytes, ``numpy.unpackbits(..., bitorder="little")`` restores them): 8x fewer D2H bytes."""
        self.packed = packed
        p = next(model.parameters())
        self.graphs = graphs
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
        if self.graphs:
            if s.graph is None or s.graph.x is not s.x or s.graph.masks is not s.masks:
                s.graph = GraphedStep(self.model, s.x, threshold=self.threshold, mean=self.mean, std=self.std,
                                      plan=s.plan, masks=s.masks)
            s.res = s.graph.replay()
            s.consumed.record(compute)
        else:
            with torch.no_grad():
                if s.x.dtype == torch.uint8:
                    logits = self.model.forward_u8(s.x, self.mean, self.std)
                else:
                    logits = self.model(s.x)
                s.consumed.record(compute)
                s.res = decode_masks(logits, threshold=self.threshold, out=s.masks, plan=s.plan)
        if self.packed:
            if getattr(s, "bits", None) is None:
                s.bits = torch.empty((self.plan.total + 7) // 8 + 1, dtype=torch.uint8, device=self.dev)
            pack_mask_bits(s.masks, self.plan.total, out=s.bits)
        s.computed.record(compute)
        with torch.cuda.stream(self.d2h):
            self.d2h.wait_event(s.computed)
            if masks_host is not None and self.packed:
                nb = (self.plan.total + 7) // 8
                masks_host[:nb].copy_(s.bits[:nb], non_blocking=True)
            elif masks_host is not None:
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


class GraphedStep:
    """Forward + decode of one fixed batch buffer captured in a CUDA graph: one host call per step instead of the
    plan's ~24 launches ("linear scaling limited only by host launch jitter -> CUDA graphs per GPU").

    ``x``: the STATIC device input (fp32 [B,C,H,W] or uint8 [B,H,W,C]); refill it in place (``x.copy_``) and call
    ``replay()``; ``result`` (DecodeResult: masks / area / bbox, and ``logits``) is rewritten by every replay.
    The programmatic-dependent-launch edges of the kernels are kept by the capture."""

    def __init__(self, model, x, boxes=None, threshold=0.5, mean=(0.5, 0.5, 0.5), std=(0.5, 0.5, 0.5),
                 plan=None, masks=None, decode=True):
        if not x.is_cuda:
            raise RuntimeError("GraphedStep needs a CUDA input buffer; there is no CPU path")
        self.model, self.x, self.threshold = model, x, threshold
        self.mean, self.std = tuple(mean), tuple(std)
        B = x.shape[0]
        H, W = (x.shape[1], x.shape[2]) if x.dtype == torch.uint8 else (x.shape[2], x.shape[3])
        self.plan = plan if plan is not None else (
            DecodePlan(identity_boxes(B, H, W) if boxes is None else boxes, x.device) if decode else None)
        self.masks = masks if masks is not None else (
            torch.empty(self.plan.total, dtype=torch.uint8, device=x.device) if decode else None)
        self.decode = decode
        self.logits = self.result = None
        side = torch.cuda.Stream(x.device)
        side.wait_stream(torch.cuda.current_stream(x.device))
        with torch.cuda.stream(side):
            for _ in range(2):      # plan creation, autotune, weight upload and occupancy queries happen here, not in the capture
                self._body()
        torch.cuda.current_stream(x.device).wait_stream(side)
        torch.cuda.synchronize(x.device)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self._body()

    def _body(self):
        with torch.no_grad():
            if self.x.dtype == torch.uint8:
                self.logits = self.model.forward_u8(self.x, self.mean, self.std)
            else:
                self.logits = self.model(self.x)
            if self.decode:
                self.result = decode_masks(self.logits, threshold=self.threshold, out=self.masks, plan=self.plan)

    def replay(self):
        self.graph.replay()
        return self.result
