"""TEST INFRASTRUCTURE ONLY -- CPU restatement of the loss of the reference's training step.

Reference: train_instance.py:299 (`criterion = torch.nn.BCELoss()`), :377-378
(`pred = model.train_batch(...)` = sigmoid of the logits, model/segment.py:531-534; `loss =
criterion(pred, mask)`; `loss.backward()`).  torch's published rules, restated in numpy fp32:
    p = 1 / (1 + exp(-x));   l = -(y * max(log p, -100) + (1 - y) * max(log(1 - p), -100));   loss = mean(l)
    dloss/dp = (p - y) / max(p * (1 - p), 1e-12) / n;     dp/dx = p * (1 - p)
Pinned in tests/test_oracle.py against torch.nn.BCELoss()(torch.sigmoid(x), y) and its autograd gradient
on CPU (the very calls the reference makes).  Only tests/ may import this file; the product never does.
"""
import numpy as np


def bce_sigmoid(logits, target):
    """-> (loss float32 scalar, dloss/dlogits float32 array of logits' shape)"""
    x = np.asarray(logits, np.float32)
    y = np.asarray(target, np.float32)
    one = np.float32(1)
    with np.errstate(over="ignore", divide="ignore"):
        p = one / (one + np.exp(-x))
        q = one - p
        lp = np.maximum(np.log(p), np.float32(-100))
        lq = np.maximum(np.log(q), np.float32(-100))
    l = -(y * lp + (one - y) * lq)
    n = x.size
    loss = np.float32(l.astype(np.float64).sum() / n)
    pq = p * q
    grad = (p - y) / np.maximum(pq, np.float32(1e-12)) * pq * np.float32(1.0 / n)
    return loss, grad.astype(np.float32)
