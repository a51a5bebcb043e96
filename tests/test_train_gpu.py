"""Training-step pieces that run in libisg (first slice): the fused sigmoid + BCE loss with its
gradient (isg_bce_sigmoid) against oracle/bce_oracle.py, and the graph-replayed step against the eager one."""
import numpy as np
import pytest
import torch

from instancesegmentation_b200 import Segment, _lib, synth
from instancesegmentation_b200.train_step import FlatBucket, GraphedTrainStep, bce_sigmoid_loss
from oracle.bce_oracle import bce_sigmoid

pytestmark = pytest.mark.gpu


def _case(shape, seed, scale=4.0):
    rng = np.random.default_rng(seed)
    x = (rng.standard_normal(shape) * scale).astype(np.float32)
    y = (rng.random(shape) > 0.7).astype(np.float32)
    return x, y


@pytest.mark.parametrize("shape", [(1,), (3, 1, 17, 29), (16, 1, 512, 512)])
def test_bce_sigmoid_fp32_matches_oracle(shape):
    x, y = _case(shape, 11)
    if x.size >= 6:
        x.reshape(-1)[:6] = [-200.0, 200.0, -90.0, 90.0, -20.0, 20.0]   # log clamp (-100) and gradient clamp (1e-12)
        y.reshape(-1)[:6] = [1, 0, 1, 0, 0, 1]
    o_loss, o_grad = bce_sigmoid(x, y)
    xt = torch.from_numpy(x).cuda().requires_grad_(True)
    loss = bce_sigmoid_loss(xt, torch.from_numpy(y).cuda())
    loss.backward()
    # fp32: device expf/logf differ from the host's by a few ulp per element; tolerance 2e-6 relative on the mean.
    # Gradient: p - y cancels where a saturated prediction is right, so an ulp of p is an absolute error of
    # ulp(1)/n there -- atol = 1e-6 of the largest possible gradient (1/n), rtol 2e-5 elsewhere.
    assert abs(loss.item() - float(o_loss)) <= 2e-6 * abs(float(o_loss))
    np.testing.assert_allclose(xt.grad.cpu().numpy(), o_grad, rtol=2e-5, atol=1e-6 / x.size)


def test_bce_sigmoid_bf16_logits_and_upstream_gradient():
    x, y = _case((4, 1, 64, 96), 12)
    xb = torch.from_numpy(x).cuda().to(torch.bfloat16)
    o_loss, o_grad = bce_sigmoid(xb.float().cpu().numpy(), y)     # the oracle sees the same rounded logits
    xt = xb.clone().requires_grad_(True)
    loss = bce_sigmoid_loss(xt, torch.from_numpy(y).cuda())
    (loss * 3.0).backward()
    assert loss.dtype == torch.float32 and xt.grad.dtype == torch.bfloat16
    assert abs(loss.item() - float(o_loss)) <= 2e-6 * abs(float(o_loss))
    # gradient is stored in bf16 (8 mantissa bits), then scaled by the upstream 3.0 in bf16
    np.testing.assert_allclose(xt.grad.float().cpu().numpy(), 3.0 * o_grad, rtol=2 ** -7, atol=1e-12)


def test_bce_sigmoid_is_deterministic_and_reusable():
    x, y = _case((8, 1, 256, 256), 13)
    xt, yt = torch.from_numpy(x).cuda(), torch.from_numpy(y).cuda()
    vals = [bce_sigmoid_loss(xt, yt).item() for _ in range(5)]     # same workspace, ticket reset by the kernel
    assert len(set(vals)) == 1
    ref = torch.nn.BCELoss()(torch.sigmoid(xt), yt).item()
    assert abs(vals[0] - ref) <= 2e-6 * abs(ref)


def test_bce_sigmoid_argument_errors():
    xt = torch.zeros(4, device="cuda")
    with pytest.raises(ValueError):
        bce_sigmoid_loss(xt, torch.zeros(5, device="cuda"))
    with pytest.raises(ValueError):
        bce_sigmoid_loss(xt[:0], xt[:0])
    with pytest.raises(TypeError):
        bce_sigmoid_loss(xt.half(), xt)
    with pytest.raises(_lib.IsgError):
        bce_sigmoid_loss(xt.cpu(), xt.cpu())
    l = _lib.lib()
    assert l.isg_bce_sigmoid(None, 0, None, 4, None, None, None, None) != 0
    assert b"null" in l.isg_last_error()


def test_graphed_step_equals_eager_step():
    """one step of GraphedTrainStep (graph replay of forward + isg_bce_sigmoid + backward) leaves the same
    loss, gradients and BatchNorm statistics as the reference's eager sequence (train_instance.py:372-378)"""
    B, S = 2, 128
    x = torch.from_numpy(synth.batch(B, S, S, seed=21)).cuda()
    y = (torch.from_numpy(synth.logits(B, S, S, seed=22)) > 0.8).float()[:, None].cuda()
    torch.manual_seed(3)
    a = Segment(3).cuda().train()
    torch.manual_seed(3)
    b = Segment(3).cuda().train()
    ba, bb = FlatBucket(a, grads_as_views=True), FlatBucket(b, grads_as_views=True)
    ba.zero_grads()
    loss_a = torch.nn.BCELoss()(a.train_batch(x), y)
    loss_a.backward()
    step = GraphedTrainStep(b, bb, x, y, autocast_dtype=None)
    loss_b = step(x, y)
    torch.cuda.synchronize()
    assert abs(loss_a.item() - loss_b.item()) <= 1e-5 * abs(loss_a.item())
    ga, gb = ba.grad, bb.grad
    assert float((ga - gb).abs().max()) <= 2e-3 * float(ga.abs().max())
    for (ka, va), (kb, vb) in zip(a.named_buffers(), b.named_buffers()):
        assert ka == kb
        if va.dtype.is_floating_point:
            assert torch.allclose(va, vb, rtol=1e-4, atol=1e-6), ka      # warm-up passes did not leak into the stats
        else:
            assert torch.equal(va, vb), ka                               # num_batches_tracked == 1 on both
    assert bb.step_count == 1
    # second replay with another batch follows the eager model too
    x2 = torch.from_numpy(synth.batch(B, S, S, seed=23)).cuda()
    ba.adam_step(ba.allreduce())
    a.invalidate()
    ba.zero_grads()
    loss_a2 = torch.nn.BCELoss()(a.train_batch(x2), y)
    loss_b2 = step(x2, y)
    torch.cuda.synchronize()
    assert abs(loss_a2.item() - loss_b2.item()) <= 5e-3 * abs(loss_a2.item())


def _bn_reference(x, gamma, beta, dy, eps=1e-5, momentum=0.1):
    """the reference's op (nn.BatchNorm2d in train mode, segment.py:41) on CPU in float64"""
    xd = x.double().cpu().requires_grad_(True)
    g = gamma.double().cpu().requires_grad_(True)
    b = beta.double().cpu().requires_grad_(True)
    rm, rv = torch.zeros(x.shape[1], dtype=torch.float64), torch.ones(x.shape[1], dtype=torch.float64)
    y = torch.nn.functional.batch_norm(xd, rm, rv, g, b, True, momentum, eps)
    y.backward(dy.double().cpu())
    return y.detach(), xd.grad, g.grad, b.grad, rm, rv


@pytest.mark.parametrize("shape", [(16, 16, 64, 64), (2, 48, 16, 16), (3, 19, 7, 9), (1, 4, 1, 2), (4, 128, 32, 32)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("two_pass", [False, True])
def test_bn_train_matches_torch_float64(shape, dtype, two_pass, monkeypatch):
    """two_pass=False: one launch per op, a thread-block cluster per channel (partials through DSMEM);
    True: the two-launch kernels with global partials (ISG_BN_TWO_PASS=1)"""
    from instancesegmentation_b200.bn import BatchNorm2d
    if two_pass:
        monkeypatch.setenv("ISG_BN_TWO_PASS", "1")
    else:
        monkeypatch.delenv("ISG_BN_TWO_PASS", raising=False)
    torch.manual_seed(sum(shape))
    C = shape[1]
    # per-channel means up to 12 sigma away from zero (the shifted sums must not cancel); the 2-element case keeps small
    # means: there x - mean is ~1e-4 of x and fp32 storage of x itself is the limit, whatever the kernel does
    spread = 30.0 if shape[0] * shape[2] * shape[3] >= 64 else 1.0
    x = (torch.randn(shape) * 2.5 + torch.linspace(-spread, spread, C).view(1, C, 1, 1)).to(dtype)
    dy = torch.randn(shape).to(dtype)
    gamma, beta = torch.rand(C) + 0.5, torch.randn(C)
    y_ref, dx_ref, dg_ref, db_ref, rm_ref, rv_ref = _bn_reference(x, gamma, beta, dy)
    m = BatchNorm2d(C).cuda().train()
    with torch.no_grad():
        m.weight.copy_(gamma)
        m.bias.copy_(beta)
    xc = x.cuda().requires_grad_(True)
    y = m(xc)
    y.backward(dy.cuda())
    y = y.detach()
    assert y.dtype == dtype and xc.grad.dtype == dtype and int(m.num_batches_tracked) == 1
    # fp32: statistics in double, elementwise math in fp32 -> 1e-5 relative to the tensor's scale;
    # bf16: the outputs are rounded to 8 mantissa bits
    tol = 1e-5 if dtype == torch.float32 else 2 ** -7
    if shape[0] * shape[2] * shape[3] < 16 and dtype == torch.float32:
        tol = 5e-3    # two elements per channel: dx is a difference of nearly equal terms scaled by invstd (37 here)

    def close(a, ref, scale_tol=tol):
        a, ref = a.double().cpu(), ref.double()
        assert float((a - ref).abs().max()) <= scale_tol * max(float(ref.abs().max()), 1e-3), float((a - ref).abs().max())

    close(y, y_ref)
    close(xc.grad, dx_ref)
    close(m.weight.grad, dg_ref, 1e-4 if dtype == torch.float32 else tol)
    close(m.bias.grad, db_ref, 1e-4 if dtype == torch.float32 else tol)
    close(m.running_mean, rm_ref, 1e-5)
    close(m.running_var, rv_ref, 1e-5)


def test_bn_train_is_deterministic_and_falls_back_outside_its_domain():
    from instancesegmentation_b200.bn import BatchNorm2d
    m = BatchNorm2d(48).cuda().train()
    x = torch.randn(8, 48, 64, 64, device="cuda")
    outs = []
    for _ in range(3):
        m.reset_running_stats()
        outs.append((m(x).clone(), m.running_var.clone()))
    assert all(torch.equal(outs[0][0], o[0]) and torch.equal(outs[0][1], o[1]) for o in outs[1:])
    m.eval()
    assert torch.allclose(m(x), torch.nn.functional.batch_norm(x, m.running_mean, m.running_var, m.weight, m.bias, False, 0.1, 1e-5))
    cpu = BatchNorm2d(3).train()
    assert cpu(torch.randn(2, 3, 4, 4)).shape == (2, 3, 4, 4)      # CPU tensors: parent class (boundary tests use it)
    l = _lib.lib()
    assert l.isg_bn_train_forward(None, 0, 1, 1, 1, None, None, None, None, None, 0.1, 1e-5, 0, None, 0, None, None,
                                  None, None, None) != 0
    assert b"null" in l.isg_last_error()


@pytest.mark.parametrize("act", ["prelu", "relu", "none"])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_conv_bn_act_block_matches_the_literal_sequence(act, dtype):
    """the reference's Conv block act(bn(conv(x))) (segment.py:34-48) in train mode: the native path (conv without
    its bias, BatchNorm + activation fused, forward and backward) against the same modules evaluated literally in
    float64 on CPU -- output, input gradient, every parameter gradient, running statistics"""
    import copy
    from instancesegmentation_b200.segment import _Unit
    torch.manual_seed(7)
    u = _Unit(12, 24, 3, act=act if act != "none" else None)
    with torch.no_grad():
        u.conv.weight.mul_(4.0)              # conv output std ~1, so bf16 rounding of it is small against the bias
        u.conv.bias.normal_(0, 2.0)          # a bias the normalisation has to cancel
        u.bn.weight.uniform_(0.5, 1.5)
        u.bn.bias.normal_()
        if act == "prelu":
            u.act.weight.uniform_(0.05, 0.5)
    ref = copy.deepcopy(u).double().train()
    x = torch.randn(4, 12, 24, 40)
    dy = torch.randn(4, 24, 24, 40)
    g = u.cuda().train()
    xg = x.cuda().requires_grad_(True)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False      # the fp32 case compares at 2e-5: no TF32 in the library conv
    if dtype == torch.bfloat16:
        with torch.autocast("cuda", dtype=torch.bfloat16):
            yg = g(xg)
    else:
        yg = g(xg)
    assert yg.dtype == dtype
    yg.backward(dy.cuda().to(dtype))
    torch.backends.cudnn.allow_tf32 = tf32
    # literal sequence in float64.  The activation's branch (z > 0) is taken from the device result: rounding moves a few
    # z across 0 (|z| tiny there, so y still agrees) and the gradient of those elements follows the branch actually taken
    xr = x.double().requires_grad_(True)
    z = ref.bn(ref.conv(xr))
    if act == "none":
        yr = z
    else:
        slope = ref.act.weight.view(1, -1, 1, 1) if act == "prelu" else 0.0
        yr = torch.where(yg.detach().cpu() > 0, z, z * slope)
    yr.backward(dy.double())
    tol = 2e-5 if dtype == torch.float32 else 3e-2     # bf16: conv inputs, weights and activations rounded to 8 bits

    def close(a, r, t=tol):
        a, r = a.double().cpu(), r.double()
        assert float((a - r).abs().max()) <= t * max(float(r.abs().max()), 1e-3), float((a - r).abs().max())

    close(yg, yr.detach())
    close(xg.grad, xr.grad)
    close(g.conv.weight.grad, ref.conv.weight.grad, tol * 5)
    close(g.bn.weight.grad, ref.bn.weight.grad, tol * 5)
    close(g.bn.bias.grad, ref.bn.bias.grad, tol * 5)
    if act == "prelu":
        close(g.act.weight.grad, ref.act.weight.grad, tol * 5)
    # the conv bias: analytically zero gradient (the literal sequence produces rounding noise there), not computed
    assert g.conv.bias.grad is None and float(ref.conv.bias.grad.abs().max()) < 1e-9
    close(g.bn.running_mean, ref.bn.running_mean, 1e-5 if dtype == torch.float32 else 1e-2)
    close(g.bn.running_var, ref.bn.running_var, 1e-5 if dtype == torch.float32 else 2e-2)


@pytest.mark.parametrize("k,d", [((3, 3), 1), ((3, 3), 2), ((3, 3), 4), ((5, 1), 1), ((1, 5), 1), ((5, 5), 1)])
@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
def test_depthwise_conv_matches_torch_float64(k, d, dtype):
    """the reference's depthwise convs (segment.py:58, :93-96: groups == channels; 3x3 dil 1/2/4, (5,1), (1,5)) in train
    mode: forward, input gradient, weight and bias gradient against F.conv2d in float64 on CPU; plane sizes that are
    not multiples of the 16x64 tile, several tiles per plane"""
    from instancesegmentation_b200.dwconv import DepthwiseConv2d
    torch.manual_seed(k[0] * 10 + k[1] + d)
    C, B, H, W = (12, 3, 37, 70) if k != (5, 1) else (48, 2, 8, 8)      # (5,1): the 8x8 planes of bottle2_x at 128x128
    pad = (d * (k[0] - 1) // 2, d * (k[1] - 1) // 2)
    m = DepthwiseConv2d(C, C, k, padding=pad, dilation=d, groups=C).cuda().train()
    x = torch.randn(B, C, H, W)
    dy = torch.randn(B, C, H, W)
    if dtype == torch.bfloat16:
        x, dy = x.to(dtype).float(), dy.to(dtype).float()     # the reference sees the same rounded inputs
    xr = x.double().requires_grad_(True)
    wr = m.weight.detach().double().cpu().requires_grad_(True)
    br = m.bias.detach().double().cpu().requires_grad_(True)
    yr = torch.nn.functional.conv2d(xr, wr, br, 1, pad, d, C)
    yr.backward(dy.double())
    xg = x.cuda().to(dtype).requires_grad_(True)
    assert m.native(xg)
    yg = m(xg)
    yg.backward(dy.cuda().to(dtype))
    assert yg.dtype == dtype and m.weight.grad.dtype == torch.float32
    tol = 1e-5 if dtype == torch.float32 else 2 ** -7       # bf16: outputs rounded to 8 bits; sums accumulate in fp32

    def close(a, r, t):
        a, r = a.double().cpu(), r.double()
        assert float((a - r).abs().max()) <= t * max(float(r.abs().max()), 1e-3), float((a - r).abs().max())

    close(yg.detach(), yr.detach(), tol)
    close(xg.grad, xr.grad, tol)
    close(m.weight.grad, wr.grad, 1e-5)      # fp32 accumulation of exactly representable products
    close(m.bias.grad, br.grad, 1e-5)
    # outside its domain the parent class runs: eval mode, stride 2
    m.eval()
    assert not m.native(xg)
    with torch.autocast("cuda", dtype=torch.bfloat16, enabled=dtype == torch.bfloat16):
        assert torch.allclose(m(xg.detach()).float(), yg.detach().float(), atol=5e-2 if dtype == torch.bfloat16 else 1e-4)
    assert not DepthwiseConv2d(C, C, 3, stride=2, padding=1, groups=C).cuda().train().native(xg)


@pytest.mark.parametrize("dtype", [torch.float32, torch.bfloat16])
@pytest.mark.parametrize("geom", [(16, 4, 8, 4, 2), (6, 8, 4, 2, 1), (3, 2, 2, 1, 0)])
def test_conv_transpose_k2s_matches_torch_float64(geom, dtype):
    """bottle6_1 (segment.py:435: ConvTranspose2d(16, 4, 8, stride 4, padding 2)) and two other kernel == 2*stride shapes:
    native forward against F.conv_transpose2d in float64; gradients (library backward of the same op) as well"""
    from instancesegmentation_b200.dwconv import ConvTranspose2dK2S
    cin, cout, k, s, p = geom
    torch.manual_seed(cin + k)
    m = ConvTranspose2dK2S(cin, cout, k, stride=s, padding=p).cuda().train()
    x = torch.randn(2, cin, 9, 13)
    if dtype == torch.bfloat16:
        x = x.to(dtype).float()
    xr = x.double().requires_grad_(True)
    wr = m.weight.detach().double().cpu().requires_grad_(True)
    br = m.bias.detach().double().cpu().requires_grad_(True)
    yr = torch.nn.functional.conv_transpose2d(xr, wr, br, s, p)
    dy = torch.randn_like(yr).float()
    if dtype == torch.bfloat16:
        dy = dy.to(dtype).float()
    yr.backward(dy.double())
    xg = x.cuda().to(dtype).requires_grad_(True)
    assert m.native(xg)
    tf32 = torch.backends.cudnn.allow_tf32
    torch.backends.cudnn.allow_tf32 = False
    yg = m(xg)
    yg.backward(dy.cuda().to(dtype))
    torch.backends.cudnn.allow_tf32 = tf32
    assert yg.shape == yr.shape and yg.dtype == dtype and m.weight.grad.dtype == torch.float32

    def close(a, r, t):
        a, r = a.double().cpu(), r.double()
        assert float((a - r).abs().max()) <= t * max(float(r.abs().max()), 1e-3), float((a - r).abs().max())

    close(yg.detach(), yr.detach(), 1e-5 if dtype == torch.float32 else 2 ** -7)
    lib_tol = 1e-4 if dtype == torch.float32 else 3e-2       # library backward; bf16: weights rounded to 8 bits there
    close(xg.grad, xr.grad, lib_tol)
    close(m.weight.grad, wr.grad, lib_tol)
    close(m.bias.grad, br.grad, lib_tol)
    m.eval()
    assert not m.native(xg)


def test_bce_sigmoid_backward_twice_with_retain_graph():
    """the gradient is a saved tensor: a second backward (retain_graph) returns it again instead of None"""
    x = torch.randn(4, 1, 32, 32, device="cuda", requires_grad=True)
    y = (torch.rand(4, 1, 32, 32, device="cuda") < 0.4).float()
    loss = bce_sigmoid_loss(x, y)
    (g1,) = torch.autograd.grad(loss, x, retain_graph=True)
    (g2,) = torch.autograd.grad(loss, x)
    assert torch.equal(g1, g2) and float(g1.abs().sum()) > 0
