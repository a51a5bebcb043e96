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

    def check_aliasing(self):
        """The parameters must still be views of ``flat`` (``model.to()`` / ``.float()`` after the bucket was built
        re-allocates them: Adam would then update memory the model no longer reads)."""
        o = 0
        for p, n in zip(self.params, self.sizes):
            if p.data_ptr() != self.flat.data_ptr() + 4 * o:
                raise RuntimeError("FlatBucket: a parameter no longer aliases the flat buffer (the model was moved or "
                                   "cast after the bucket was built); build a new FlatBucket")
            o += n

    def state_dict(self):
        """Optimizer state in ``torch.optim.Adam.state_dict()`` format (per-parameter ``step`` / ``exp_avg`` /
        ``exp_avg_sq``, one param group), as the reference saves it next to the model (train_instance.py:497-503)."""
        state, o = {}, 0
        for i, (p, n) in enumerate(zip(self.params, self.sizes)):
            state[i] = {"step": torch.tensor(float(self.step_count)),
                        "exp_avg": self.exp_avg[o:o + n].view_as(p).clone(),
                        "exp_avg_sq": self.exp_avg_sq[o:o + n].view_as(p).clone()}
            o += n
        group = {"lr": 1e-3, "betas": (0.9, 0.999), "eps": 1e-8, "weight_decay": 0, "amsgrad": False,
                 "params": list(range(len(self.params)))}
        return {"state": state, "param_groups": [group]}

    def load_state_dict(self, sd):
        """Resume from a ``torch.optim.Adam`` (or ``FlatBucket``) state dict: moments and step count."""
        state = sd["state"]
        steps, o = set(), 0
        for i, (p, n) in enumerate(zip(self.params, self.sizes)):
            st = state.get(i, state.get(str(i)))
            if st is None:                      # a parameter that never got a gradient has no entry in torch's dict
                self.exp_avg[o:o + n].zero_(); self.exp_avg_sq[o:o + n].zero_()
            else:
                if st["exp_avg"].numel() != n:
                    raise ValueError(f"optimizer state of parameter {i} has {st['exp_avg'].numel()} elements, expected {n}")
                self.exp_avg[o:o + n].copy_(st["exp_avg"].reshape(-1))
                self.exp_avg_sq[o:o + n].copy_(st["exp_avg_sq"].reshape(-1))
                steps.add(int(float(st["step"])))
            o += n
        if len(steps) > 1:
            raise ValueError(f"per-parameter step counts differ ({sorted(steps)}): the flat update keeps one")
        self.step_count = steps.pop() if steps else 0

    def adam_step(self, grad_scale, lr=1e-3, betas=(0.9, 0.999), eps=1e-8):
        """torch.optim.Adam defaults of train_instance.py:297, in libisg (CUDA only)."""
        if not self.flat.is_cuda:
            raise _lib.IsgError("adam_step runs in libisg on CUDA only; there is no CPU path")
        self.check_aliasing()
        self.step_count += 1
        with torch.cuda.device(self.flat.device):
            _lib.check(_lib.lib().isg_adam_step(
                self.flat.data_ptr(), self.grad.data_ptr(), self.exp_avg.data_ptr(), self.exp_avg_sq.data_ptr(),
                self.numel, float(grad_scale), float(lr), float(betas[0]), float(betas[1]), float(eps),
                self.step_count, torch.cuda.current_stream().cuda_stream))


_BCE_WS = {}


def _bce_workspace(dev):
    """per-device scratch of isg_bce_sigmoid (per-block partial sums + ticket), zeroed once"""
    key = (dev.type, dev.index)
    if key not in _BCE_WS:
        _BCE_WS[key] = torch.zeros(_lib.lib().isg_bce_workspace_bytes() // 8, dtype=torch.float64, device=dev)
    return _BCE_WS[key]


class _BceSigmoid(torch.autograd.Function):
    """loss = BCELoss()(sigmoid(logits), target) and its gradient w.r.t. the logits in one libisg launch."""

    @staticmethod
    def forward(ctx, logits, target):
        if not logits.is_cuda:
            raise _lib.IsgError("bce_sigmoid_loss runs in libisg on CUDA only; there is no CPU path")
        if logits.dtype not in (torch.float32, torch.bfloat16):
            raise TypeError(f"logits must be float32 or bfloat16, got {logits.dtype}")
        if target.shape != logits.shape:
            raise ValueError(f"target shape {tuple(target.shape)} differs from logits {tuple(logits.shape)}")
        if logits.numel() == 0:
            raise ValueError("bce_sigmoid_loss of an empty batch")
        logits = logits.contiguous()
        target = target.to(torch.float32).contiguous()
        need_grad = ctx.needs_input_grad[0]
        dlogits = torch.empty_like(logits) if need_grad else None
        loss = torch.empty((), dtype=torch.float32, device=logits.device)
        with torch.cuda.device(logits.device):
            _lib.check(_lib.lib().isg_bce_sigmoid(
                logits.data_ptr(), int(logits.dtype == torch.bfloat16), target.data_ptr(), logits.numel(),
                dlogits.data_ptr() if need_grad else None, _bce_workspace(logits.device).data_ptr(),
                loss.data_ptr(), torch.cuda.current_stream().cuda_stream))
        if need_grad:
            ctx.save_for_backward(dlogits)     # survives retain_graph / a second backward, freed with the graph
        ctx.has_grad = need_grad
        return loss

    @staticmethod
    def backward(ctx, g):
        if not ctx.has_grad:
            return None, None
        (d,) = ctx.saved_tensors
        return d * g.to(d.dtype), None


def bce_sigmoid_loss(logits, target):
    """``torch.nn.BCELoss()(torch.sigmoid(logits), target)`` (train_instance.py:299, :377-378, segment.py:534)
    computed from the raw logits of ``Segment.forward`` -- same value, same gradient, one kernel."""
    return _BceSigmoid.apply(logits, target)


class GraphedTrainStep:
    """The step of train_instance.py:372-380 with forward, loss and backward replayed from ONE CUDA graph.

    The step is ~1.5 k small library launches for 16 images; captured once, a step is: copy the batch into
    the static input buffers, replay, all-reduce the flat bucket, ``isg_adam_step``.  BatchNorm running
    statistics touched by the capture warm-up are restored, so step k of this class equals step k of
    ``train_step`` on the same data."""

    def __init__(self, model, bucket, images, masks, autocast_dtype=torch.bfloat16, warmup=4, cudnn_benchmark=True):
        """``cudnn_benchmark``: let cuDNN time its algorithms during the warm-up passes; the captured graph then holds
        the chosen kernels (10.7 vs 12.3 ms of GPU time per step at 16 x 512^2); the global flag is restored."""
        if not bucket.grads_as_views:
            raise ValueError("GraphedTrainStep needs FlatBucket(model, grads_as_views=True): the graph writes "
                             "the gradients at fixed addresses")
        if not images.is_cuda:
            raise _lib.IsgError("GraphedTrainStep runs on CUDA only")
        self.model, self.bucket, self.autocast_dtype = model, bucket, autocast_dtype
        self.x, self.y = torch.empty_like(images), torch.empty_like(masks, dtype=torch.float32)
        self.x.copy_(images)
        self.y.copy_(masks)
        model.train()
        saved = {k: v.clone() for k, v in model.named_buffers()}
        flag = torch.backends.cudnn.benchmark
        torch.backends.cudnn.benchmark = bool(cudnn_benchmark)
        side = torch.cuda.Stream(device=images.device)
        side.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(side):
            for _ in range(warmup):
                self._body()
        torch.cuda.current_stream().wait_stream(side)
        self.graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(self.graph):
            self.loss = self._body()
        torch.backends.cudnn.benchmark = flag
        with torch.no_grad():
            for k, v in model.named_buffers():
                v.copy_(saved[k])

    def _body(self):
        """One forward + loss + backward.  Gradients are produced as fresh tensors (``p.grad = None`` first, so autograd
        assigns instead of launching one accumulate kernel per parameter) and copied into the zeroed flat bucket with a
        single multi-tensor copy; afterwards every ``p.grad`` is a view of the bucket again."""
        from .bn import deferred_counters
        b = self.bucket
        b.grad.zero_()
        for p in b.params:
            p.grad = None
        with deferred_counters():
            if self.autocast_dtype is None:
                logits = self.model(self.x)
            else:
                with torch.autocast("cuda", dtype=self.autocast_dtype):
                    logits = self.model(self.x)
        loss = bce_sigmoid_loss(logits, self.y)
        loss.backward()
        dst, src, o = [], [], 0
        for p, n in zip(b.params, b.sizes):
            if p.grad is not None:
                dst.append(b.grad[o:o + n].view_as(p))
                src.append(p.grad)
            o += n
        if dst:
            torch._foreach_copy_(dst, src)
        b.bind_grads()
        return loss.detach()

    def __call__(self, images, masks):
        self.x.copy_(images, non_blocking=True)
        self.y.copy_(masks, non_blocking=True)
        self.graph.replay()
        self.bucket.adam_step(self.bucket.allreduce())
        self.model.invalidate()
        return self.loss


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
