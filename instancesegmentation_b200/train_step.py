"""Data-parallel exchange of the training step (reference train_instance.py:372-380).

One process per GPU.  Every parameter gradient is packed into ONE flat fp32 bucket (257 145
floats = 1.03 MB for ``Segment(3)``; parameters that got no gradient -- the reference has three
unused ``prelu.weight`` tensors -- are zero filled), the bucket is summed with a single
``torch.distributed.all_reduce`` (NCCL over NVLink on the box, gloo in the CPU tests) and the
Adam update (train_instance.py:297 defaults) runs in libisg over the flat views with the
1/world_size scale fused (``isg_adam_step``).  BatchNorm statistics stay per rank, exactly as
independent replicas of the reference would behave (it has no SyncBN).
"""
import torch
import torch.distributed as dist

from . import _lib


class FlatBucket:
    """Flat fp32 views of all parameters of ``model`` (in ``named_parameters`` order)."""

    def __init__(self, model, grads_as_views=False):
        """``grads_as_views``: every ``p.grad`` is a view of the flat gradient bucket, so autograd accumulates
        straight into it and ``pack_grads`` has nothing to copy (366 small copies otherwise); the bucket is
        zeroed once per step instead of setting the grads to None."""
        self.params = [p for _, p in model.named_parameters()]
        self.sizes = [p.numel() for p in self.params]
        self.numel = sum(self.sizes)
        dev = self.params[0].device
        self.grad = torch.zeros(self.numel, dtype=torch.float32, device=dev)
        # parameters become views of one flat tensor so that the optimizer is a single launch
        self.flat = torch.empty(self.numel, dtype=torch.float32, device=dev)
        o = 0
        with torch.no_grad():
            for p, n in zip(self.params, self.sizes):
                self.flat[o:o + n].copy_(p.reshape(-1))
                p.data = self.flat[o:o + n].view_as(p)
                o += n
        self.exp_avg = torch.zeros_like(self.flat)
        self.exp_avg_sq = torch.zeros_like(self.flat)
        self.step_count = 0
        self.grads_as_views = bool(grads_as_views)
        if self.grads_as_views:
            self.bind_grads()

    def bind_grads(self):
        o = 0
        for p, n in zip(self.params, self.sizes):
            p.grad = self.grad[o:o + n].view_as(p)
            o += n

    def zero_grads(self):
        """Start of a step: one memset (views) or grads to None (packed mode)."""
        if self.grads_as_views:
            self.grad.zero_()
        else:
            for p in self.params:
                p.grad = None

    def pack_grads(self):
        if self.grads_as_views:      # autograd wrote into the bucket already (unused parameters stay zero)
            return self.grad
        o = 0
        for p, n in zip(self.params, self.sizes):
            if p.grad is None:
                self.grad[o:o + n].zero_()
            else:
                self.grad[o:o + n].copy_(p.grad.reshape(-1))
            o += n
        return self.grad

    def allreduce(self):
        """Sum the bucket over all ranks (one collective); returns the 1/world scale to apply."""
        if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
            dist.all_reduce(self.grad, op=dist.ReduceOp.SUM)
            return 1.0 / dist.get_world_size()
        return 1.0

    def adam_step(self, grad_scale, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        """torch.optim.Adam defaults of train_instance.py:297, in libisg (CUDA only)."""
        if not self.flat.is_cuda:
            raise _lib.IsgError("adam_step runs in libisg on CUDA only; there is no CPU path")
        self.step_count += 1
        with torch.cuda.device(self.flat.device):
            _lib.check(_lib.lib().isg_adam_step(
                self.flat.data_ptr(), self.grad.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.numel, float(grad_scale), float(lr), float(betas[0]), float(betas[1]), float(eps),
                self.step_count, torch.cuda.current_stream().cuda_stream))


def train_step(model, bucket, images, masks, criterion=None):
    """forward -> BCE -> backward -> one all-reduce -> fused Adam (train_instance.py:372-380)."""
    criterion = criterion or torch.nn.BCELoss()
    model.train()
    bucket.zero_grads()
    loss = criterion(model.train_batch(images), masks)
    loss.backward()
    bucket.pack_grads()
    scale = bucket.allreduce()
    bucket.adam_step(scale)
    model.invalidate()
    return loss.detach()
