"""Train-mode BatchNorm(+PReLU) kernels against the HBM roofline: isg_bn_train_forward (3 passes over the tensor: read,
read, write) and isg_bn_train_backward (5 passes: dy, x, dy, x, dx) on the layer shapes of a 16-image step at 512x512,
bf16, called through the C ABI, `reps` launches captured in one CUDA graph (no host gaps), timed with CUDA events,
tensors rotating over > 126 MB so L2 does not serve the re-reads of another launch.
Usage: python scripts/bn_time.py [once]   ("once" = one forward+backward of the largest shape, for ncu)"""
import json
import sys

import torch

sys.path.insert(0, ".")
from instancesegmentation_b200 import _lib  # noqa: E402

once = len(sys.argv) > 1 and sys.argv[1] == "once"
shapes = [(16, 16, 256, 256)] if once else [(16, 16, 256, 256), (16, 16, 128, 128), (16, 48, 64, 64), (16, 128, 32, 32),
                                             (16, 4, 128, 128), (16, 48, 128, 128)]
try:
    peak = json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]
except Exception:
    peak = None
l = _lib.lib()
dev = torch.device("cuda")
for shape in shapes:
    B, C, H, W = shape
    nbytes = B * C * H * W * 2
    n_rot = 1 if once else max(2, int(400e6 // nbytes))
    xs = [torch.randn(shape, device=dev).to(torch.bfloat16) for _ in range(n_rot)]
    dys = [torch.randn(shape, device=dev).to(torch.bfloat16) for _ in range(n_rot)]
    out = [torch.empty_like(xs[0]) for _ in range(2)]
    g, b, slope = torch.ones(C, device=dev), torch.zeros(C, device=dev), torch.full((C,), 0.25, device=dev)
    rm, rv = torch.zeros(C, device=dev), torch.ones(C, device=dev)
    mean, invstd = torch.empty(C, device=dev), torch.empty(C, device=dev)
    dg, db, ds = torch.empty(C, device=dev), torch.empty(C, device=dev), torch.empty(C, device=dev)
    ws = torch.empty(C * l.isg_bn_max_split() * 3, dtype=torch.float64, device=dev)

    def fwd(i, st):
        _lib.check(l.isg_bn_train_forward(xs[i % n_rot].data_ptr(), 1, B, C, H * W, g.data_ptr(), b.data_ptr(), None,
                                          rm.data_ptr(), rv.data_ptr(), 0.1, 1e-5, 2, slope.data_ptr(), C,
                                          out[i % 2].data_ptr(), mean.data_ptr(), invstd.data_ptr(), ws.data_ptr(), st))

    def bwd(i, st):
        _lib.check(l.isg_bn_train_backward(dys[i % n_rot].data_ptr(), xs[i % n_rot].data_ptr(), 1, B, C, H * W,
                                           g.data_ptr(), b.data_ptr(), mean.data_ptr(), invstd.data_ptr(), 2,
                                           slope.data_ptr(), C, out[i % 2].data_ptr(), dg.data_ptr(), db.data_ptr(),
                                           ds.data_ptr(), ws.data_ptr(), st))

    if once:
        st = torch.cuda.current_stream().cuda_stream
        fwd(0, st)
        bwd(0, st)
        torch.cuda.synchronize()
        continue
    reps = 4 * n_rot
    res = {}
    for name, fn, passes in (("forward", fwd, 3), ("backward", bwd, 5)):
        s = torch.cuda.Stream()
        s.wait_stream(torch.cuda.current_stream())
        with torch.cuda.stream(s):
            for i in range(n_rot):
                fn(i, s.cuda_stream)
        torch.cuda.current_stream().wait_stream(s)
        torch.cuda.synchronize()
        graph = torch.cuda.CUDAGraph()
        with torch.cuda.graph(graph):
            st = torch.cuda.current_stream().cuda_stream
            for i in range(reps):
                fn(i, st)
        graph.replay()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(3):
            graph.replay()
        e1.record()
        torch.cuda.synchronize()
        us = e0.elapsed_time(e1) * 1e3 / (3 * reps)
        res[name] = (us, passes * nbytes / us / 1e3)
    frac = (lambda v: f" = {v / peak:.0%} of {peak:.0f}") if peak else (lambda v: "")
    print(f"{shape} bf16 {nbytes / 1e6:5.1f} MB: forward {res['forward'][0]:6.1f} us {res['forward'][1]:6.0f} GB/s"
          f"{frac(res['forward'][1])}   backward {res['backward'][0]:6.1f} us {res['backward'][1]:6.0f} GB/s{frac(res['backward'][1])}")
