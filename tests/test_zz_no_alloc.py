"""isg_bce_sigmoid_static: the allocation-free variant of isg_bce_sigmoid (scratch in a file-scope device array of
libisg instead of a caller workspace; no counterpart in the reference).
Called through the C ABI and checked against the existing entry point on the same inputs, twice in a row (the array
keeps the previous call's contents).  Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest
import torch

from instancesegmentation_b200 import _lib

pytestmark = pytest.mark.gpu


def _ptr(t):
    return C.c_void_p(t.data_ptr())


def _run_both(x, y):
    """(loss, grad) of the existing entry point and of the static one on logits x, targets y"""
    l = _lib.lib()
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    bf16 = int(x.dtype == torch.bfloat16)
    ws = torch.zeros(l.isg_bce_workspace_bytes() // 8, dtype=torch.float64, device=x.device)
    g0, g1 = torch.empty_like(x), torch.empty_like(x)
    l0, l1 = torch.zeros(1, device=x.device), torch.zeros(1, device=x.device)
    _lib.check(l.isg_bce_sigmoid(_ptr(x), bf16, _ptr(y), x.numel(), _ptr(g0), _ptr(ws), _ptr(l0), st))
    _lib.check(l.isg_bce_sigmoid_static(_ptr(x), bf16, _ptr(y), x.numel(), _ptr(g1), _ptr(l1), st))
    torch.cuda.synchronize()
    return (l0.item(), g0), (l1.item(), g1)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_static_scratch_variant_matches_workspace_variant_twice(dtype):
    rng = np.random.default_rng(11)
    for call, n in enumerate((3 * 512 * 512 + 7, 1000)):   # second call: smaller grid on the array the first one left behind
        x = torch.from_numpy(rng.normal(0, 3, n).astype(np.float32)).cuda().to(dtype)
        y = torch.from_numpy((rng.random(n) < 0.3).astype(np.float32)).cuda()
        (loss0, g0), (loss1, g1) = _run_both(x, y)
        # same grid and sum order: tolerance of tests/test_train_gpu.py (2e-6 relative on the loss, 2e-5 on the gradient)
        assert abs(loss1 - loss0) <= 2e-6 * abs(loss0), (call, loss0, loss1)
        scale = float(g0.float().abs().max())
        assert float((g1.float() - g0.float()).abs().max()) <= 2e-5 * scale, call
        assert np.isfinite(loss1) and loss1 > 0


def test_static_scratch_variant_allocates_nothing_and_checks_its_limit():
    l = _lib.lib()
    n = 4096
    x = torch.randn(n, device="cuda")
    y = torch.zeros(n, device="cuda")
    g, loss = torch.empty_like(x), torch.zeros(1, device="cuda")
    st = C.c_void_p(torch.cuda.current_stream().cuda_stream)
    _lib.check(l.isg_bce_sigmoid_static(_ptr(x), 0, _ptr(y), n, _ptr(g), _ptr(loss), st))   # set-up call of this device
    torch.cuda.synchronize()
    free0 = torch.cuda.mem_get_info()[0]
    for _ in range(3):
        _lib.check(l.isg_bce_sigmoid_static(_ptr(x), 0, _ptr(y), n, _ptr(g), _ptr(loss), st))
    torch.cuda.synchronize()
    assert torch.cuda.mem_get_info()[0] == free0            # no device allocation by the calls
    limit = l.isg_bce_sigmoid_static_max_n()
    assert limit == 1 << 26
    assert l.isg_bce_sigmoid_static(_ptr(x), 0, _ptr(y), limit + 1, _ptr(g), _ptr(loss), st) == -1
    assert b"static-scratch limit" in l.isg_last_error()
    assert l.isg_bce_sigmoid_static(None, 0, _ptr(y), n, _ptr(g), _ptr(loss), st) == -1
    assert l.isg_bce_sigmoid_static(_ptr(x), 0, _ptr(y), 0, _ptr(g), _ptr(loss), st) == -1
