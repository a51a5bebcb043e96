"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the reference forward.

Restates ``Segment.forward`` (reference model/segment.py:466-508) as a pure
function of a ``state_dict`` using torch CPU fp32 ops in eval mode (BatchNorm
with running statistics).  It exists so the parity tests can run on the GPU box,
where /root/reference is absent.  Pinned by tests/test_oracle.py against (a) the
unmodified reference module imported from /root/reference (when present) and
(b) the committed fixtures in tests/golden/ written by make_golden.py from that
reference.  Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
--impl reference legs may import this file; the product never does.
"""
import torch
import torch.nn.functional as F

EPS = 1e-5  # nn.BatchNorm2d default (segment.py:41)


def _bn(sd, p, x):
    return F.batch_norm(x, sd[p + ".running_mean"], sd[p + ".running_var"],
                        sd[p + ".weight"], sd[p + ".bias"], False, 0.0, EPS)


def _unit(sd, p, x, act, **kw):
    """Conv wrapper, segment.py:34-45: act(bn(conv(x)))."""
    y = _bn(sd, p + ".bn", F.conv2d(x, sd[p + ".conv.weight"], sd[p + ".conv.bias"], **kw))
    if act == "prelu":
        return F.prelu(y, sd[p + ".act.weight"])
    return F.relu(y) if act == "relu" else y


def _res3(sd, p, x, d):
    """Bottleneck3x3, segment.py:52-79."""
    m = sd[p + ".convs.1.conv.weight"].shape[0]
    y = _unit(sd, p + ".convs.0", x, "prelu")
    y = _unit(sd, p + ".convs.1", y, "prelu", padding=d, dilation=d, groups=m)
    y = _unit(sd, p + ".convs.2", y, None)
    return F.prelu(y + x, sd[p + ".prelu.weight"])


def _res5(sd, p, x):
    """Bottleneck5x5, segment.py:82-111 (bare (5,1) depthwise, then (1,5) unit)."""
    m = sd[p + ".convs.1.weight"].shape[0]
    y = _unit(sd, p + ".convs.0", x, "prelu")
    y = F.conv2d(y, sd[p + ".convs.1.weight"], sd[p + ".convs.1.bias"], padding=(2, 0), groups=m)
    y = _unit(sd, p + ".convs.2", y, "prelu", padding=(0, 2), groups=m)
    y = _unit(sd, p + ".convs.3", y, None)
    return F.prelu(y + x, sd[p + ".prelu.weight"])


def _down(sd, p, x):
    """BottleneckDown2, segment.py:114-150; returns (out, max-pooled input)."""
    m = sd[p + ".convs.1.conv.weight"].shape[0]
    y = _unit(sd, p + ".convs.0", x, "prelu", stride=2)
    y = _unit(sd, p + ".convs.1", y, "prelu", padding=1, groups=m)
    y = _unit(sd, p + ".convs.2", y, None)
    pooled = F.max_pool2d(x, 2, 2)
    return F.prelu(y + _unit(sd, p + ".convm.0", pooled, None), sd[p + ".prelu.weight"]), pooled


def _proj(sd, p, x, last):
    """BottleneckDim_Res, segment.py:153-209 (PReLU inside both variants)."""
    m = sd[p + ".convs.1.conv.weight"].shape[0]
    y = _unit(sd, p + ".convs.0", x, "prelu")
    y = _unit(sd, p + ".convs.1", y, "prelu", padding=1, groups=m)
    y = _unit(sd, p + ".convs.2", y, None) + _unit(sd, p + ".resconv.0", x, None)
    return F.prelu(y, sd[p + ".prelu.weight"]) if last == "prelu" else F.relu(y)


def _plain(sd, p, x):
    """BottleneckDim(usePrelu=False), segment.py:232-261 (dense 3x3, ReLU)."""
    y = _unit(sd, p + ".convs.0", x, "relu")
    y = _unit(sd, p + ".convs.1", y, "relu", padding=1)
    return F.relu(_unit(sd, p + ".convs.2", y, None) + x)


def _up(sd, p, x, skip):
    """BottleneckUp_Res(_Other), segment.py:296-344."""
    y = _unit(sd, p + ".convs.0", x, "relu")
    y = F.conv_transpose2d(y, sd[p + ".convs.1.weight"], sd[p + ".convs.1.bias"], stride=2, padding=1)
    y = F.relu(_bn(sd, p + ".convs.2", y))
    y = _unit(sd, p + ".convs.4", y, None)
    side = torch.cat((_unit(sd, p + ".conv2.0", x, None), skip), 1)
    side = F.interpolate(side, scale_factor=2, mode="nearest")
    side = F.conv2d(side, sd[p + ".uppool.1.weight"], sd[p + ".uppool.1.bias"])
    return F.relu(y + side)


_DILS = (1, 2, 1, 4)


def _wide(sd, p, x, taps=None):
    for i, d in enumerate(_DILS):
        x = _res3(sd, f"{p}.{i}", x, d)
        if taps is not None:
            taps[f"{p}.{i}"] = x
    return _res5(sd, f"{p}.4", x)


@torch.no_grad()
def segment_forward(sd, x, taps=None, dtype=torch.float32):
    """Raw logits [B,1,H,W] for CPU ``x`` [B,C,H,W]; segment.py:466-508.

    ``taps`` (optional dict) receives the named intermediate activations used by
    the layer-wise parity tests.  ``dtype=torch.float64`` evaluates the same graph in
    double precision: the yardstick for how far the reference's own fp32 arithmetic is
    from the exact value of its graph.
    """
    sd = {k: (v.detach().to("cpu", dtype) if v.dtype.is_floating_point else v.detach().to("cpu"))
          for k, v in sd.items()}
    x = x.detach().to("cpu", dtype)
    t = {} if taps is None else taps
    y = _unit(sd, "init_conv.layer1", x, "prelu", stride=2, padding=2)
    t["stem1"] = y
    y = _unit(sd, "init_conv.layer2", y, "prelu", stride=2, padding=2)
    t["init_down"] = s = torch.cat((F.max_pool2d(x, 4, 4), y), 1)          # :28-31
    t["bottle1_down"], t["keep1"] = d1, keep1 = _down(sd, "bottle1_1", s)  # :478
    y = d1
    for i in range(4):
        y = _res3(sd, f"bottle1_x.{i}", y, 1)                               # :479
        t[f"bottle1_x.{i}"] = y
    t["bottle1_5"] = y
    t["bottle2_down"], t["keep2"] = d2, keep2 = _down(sd, "bottle2_1", y)  # :482
    t["bottle2_8"] = y = _wide(sd, "bottle2_x", d2, t)                      # :483 (per-block taps bottle2_x.0..3; .4 = bottle2_8)
    t["bottle3_1"] = y = _proj(sd, "bottle3_1", torch.cat((y, d2), 1), "prelu")  # :485-488
    t["bottle3_8"] = y = _wide(sd, "bottle3_x", y, t)                       # :489
    t["bottle4_1"] = y = _up(sd, "bottle4_1up", y, keep2)                   # :492
    t["bottle4_2"] = y = _proj(sd, "bottle4_2", torch.cat((d1, y), 1), "relu")   # :494-496
    t["bottle4_3"] = y = _plain(sd, "bottle4_3", y)                         # :497
    t["bottle5_1"] = y = _up(sd, "bottle5_1up", y, keep1)                   # :500
    t["bottle5_2"] = y = _plain(sd, "bottle5_2", y)                         # :501
    t["bottle6_1"] = y = F.conv_transpose2d(y, sd["bottle6_1.weight"], sd["bottle6_1.bias"],
                                            stride=4, padding=2)            # :504
    t["logits"] = y = F.conv2d(y, sd["bottle6_2.weight"], sd["bottle6_2.bias"], padding=1)  # :505
    return y
