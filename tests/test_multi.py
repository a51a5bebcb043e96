"""Multi-process host logic on CPU (gloo, world size 2): the flat gradient bucket of the training
step and the contiguous batch shard of inference.  No kernels run here."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from instancesegmentation_b200 import Segment, synth
from instancesegmentation_b200.train_step import FlatBucket


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


def shard(n_items, rank, world):
    """contiguous split used by bench.py / infer: rank r owns [lo, hi)"""
    per = (n_items + world - 1) // world
    return min(rank * per, n_items), min((rank + 1) * per, n_items)


def test_shard_partition_is_exact():
    for n in (0, 1, 7, 64, 512, 513):
        for world in (1, 2, 4, 8):
            spans = [shard(n, r, world) for r in range(world)]
            covered = [i for lo, hi in spans for i in range(lo, hi)]
            assert covered == list(range(n))


def _worker(rank, world, port, out):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    torch.manual_seed(0)                      # identical replicas
    model = Segment(3)
    bucket = FlatBucket(model)
    x = torch.from_numpy(synth.batch(4, 32, 32, seed=5))
    y = (torch.from_numpy(synth.logits(4, 32, 32, seed=6)) > 0).float()[:, None]
    lo, hi = shard(4, rank, world)
    model.train()
    loss = torch.nn.BCELoss()(model.train_batch(x[lo:hi]), y[lo:hi])
    loss.backward()
    bucket.pack_grads()
    scale = bucket.allreduce()
    if rank == 0:
        out["grad"] = (bucket.grad * scale).clone().numpy()
        out["numel"] = bucket.numel
        out["unused"] = [n for n, p in model.named_parameters() if p.grad is None]
    dist.destroy_process_group()


def test_bucket_allreduce_matches_two_half_batches():
    world, port = 2, _free_port()
    with mp.Manager() as mgr:
        out = mgr.dict()
        mp.spawn(_worker, args=(world, port, out), nprocs=world, join=True)
        got, numel, unused = out["grad"], out["numel"], list(out["unused"])
    assert numel == 257145                       # parameters of Segment(3)
    assert sorted(unused) == ["bottle4_2.prelu.weight", "bottle4_3.prelu.weight", "bottle5_2.prelu.weight"]
    # single-process reference: mean of the two half-batch gradients (BatchNorm statistics are
    # per rank in the reference -- it has no SyncBN -- so halves are evaluated separately)
    torch.manual_seed(0)
    model = Segment(3)
    x = torch.from_numpy(synth.batch(4, 32, 32, seed=5))
    y = (torch.from_numpy(synth.logits(4, 32, 32, seed=6)) > 0).float()[:, None]
    acc = None
    for r in range(world):
        lo, hi = shard(4, r, world)
        model.zero_grad(set_to_none=True)
        model.train()
        torch.nn.BCELoss()(model.train_batch(x[lo:hi]), y[lo:hi]).backward()
        flat = torch.cat([(p.grad if p.grad is not None else torch.zeros_like(p)).reshape(-1)
                          for _, p in model.named_parameters()])
        acc = flat if acc is None else acc + flat
    want = (acc / world).numpy()
    np.testing.assert_allclose(got, want, rtol=1e-5, atol=1e-7)


def test_flat_views_keep_state_dict_keys():
    model = Segment(3)
    keys = list(model.state_dict().keys())
    bucket = FlatBucket(model)
    assert list(model.state_dict().keys()) == keys
    with torch.no_grad():
        bucket.flat.add_(1.0)                     # parameters are views of the flat tensor
    assert torch.allclose(next(iter(model.parameters())).reshape(-1)[:4], bucket.flat[:4])


def test_grads_as_bucket_views_equal_packed_grads():
    """gradient-as-bucket-view mode: autograd accumulates into the flat bucket; same numbers as packing"""
    torch.manual_seed(3)
    a, b = Segment(3), Segment(3)
    b.load_state_dict(a.state_dict())
    ba, bb = FlatBucket(a), FlatBucket(b, grads_as_views=True)
    x = torch.from_numpy(synth.batch(2, 32, 32, seed=5))
    y = (torch.from_numpy(synth.logits(2, 32, 32, seed=6)) > 0).float()[:, None]
    for step in range(2):            # the second step checks that zero_grads really clears the views
        for m, bk in ((a, ba), (b, bb)):
            m.train()
            bk.zero_grads()
            torch.nn.BCELoss()(m.train_batch(x), y).backward()
            bk.pack_grads()
        assert torch.equal(ba.grad, bb.grad)
        assert all(p.grad.data_ptr() >= bb.grad.data_ptr() for p in bb.params)


@pytest.mark.gpu
def test_adam_step_matches_torch_adam():
    torch.manual_seed(1)
    model = Segment(3).cuda()
    ref = Segment(3).cuda()
    ref.load_state_dict(model.state_dict())
    bucket = FlatBucket(model)
    opt = torch.optim.Adam(ref.parameters(), lr=1e-3)
    x = torch.from_numpy(synth.batch(2, 32, 32, seed=5)).cuda()
    y = (torch.from_numpy(synth.logits(2, 32, 32, seed=6)) > 0).float()[:, None].cuda()
    for _ in range(3):
        # one backward (on the torch replica); both optimizers see the SAME gradients, so the comparison is
        # of the update rule alone (two backward passes differ in the last bits and Adam's first steps are
        # sign-like, which would amplify that to ~lr)
        ref.train()
        ref.zero_grad(set_to_none=True)
        torch.nn.BCELoss()(ref.train_batch(x), y).backward()
        for pm, pr in zip(model.parameters(), ref.parameters()):
            pm.grad = None if pr.grad is None else pr.grad.clone()
        bucket.pack_grads()
        bucket.adam_step(bucket.allreduce())
        opt.step()
        with torch.no_grad():   # keep the replicas' inputs to the next backward identical
            for pm, pr in zip(model.parameters(), ref.parameters()):
                if pr.grad is not None:
                    assert torch.allclose(pm, pr, rtol=1e-5, atol=1e-7)
    for (n, a), (_, b) in zip(model.named_parameters(), ref.named_parameters()):
        if b.grad is None:
            continue
        assert torch.allclose(a, b, rtol=1e-4, atol=1e-6), n


def test_flat_bucket_state_dict_round_trips_with_torch_adam_and_detects_broken_aliasing():
    """FlatBucket.state_dict() has torch.optim.Adam's format (the reference checkpoints optimizer.state_dict() next to
    the model, train_instance.py:497-503) and load_state_dict() accepts a real Adam state; moving / casting the model
    after the bucket was built is detected instead of silently training stale tensors."""
    import torch
    from instancesegmentation_b200 import Segment
    from instancesegmentation_b200.train_step import FlatBucket
    torch.manual_seed(0)
    m = Segment(3)
    opt = torch.optim.Adam(m.parameters())
    for p in m.parameters():
        p.grad = torch.randn_like(p) * 1e-2
    opt.step(); opt.step()
    ref = opt.state_dict()
    b = FlatBucket(m)
    b.load_state_dict(ref)
    assert b.step_count == 2
    sd = b.state_dict()
    assert set(sd) == {"state", "param_groups"} and sd["param_groups"][0]["params"] == list(range(len(b.params)))
    for i, st in ref["state"].items():
        assert torch.equal(sd["state"][i]["exp_avg"], st["exp_avg"]) and torch.equal(sd["state"][i]["exp_avg_sq"], st["exp_avg_sq"])
        assert float(sd["state"][i]["step"]) == float(st["step"])
    opt2 = torch.optim.Adam(m.parameters())
    opt2.load_state_dict(sd)                       # torch accepts it back
    b2 = FlatBucket(Segment(3))
    b2.load_state_dict(sd)
    assert torch.equal(b2.exp_avg, b.exp_avg) and torch.equal(b2.exp_avg_sq, b.exp_avg_sq) and b2.step_count == 2
    b.check_aliasing()
    m.double()                                     # re-allocates every parameter
    try:
        b.check_aliasing()
        raise AssertionError("broken aliasing was not detected")
    except RuntimeError as e:
        assert "no longer aliases" in str(e)
