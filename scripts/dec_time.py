"""Decode time at cfg2 (64 instances, identity 512x512): CUDA events over 50 launches, inputs rotated (> L2)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import synth
from instancesegmentation_b200.decode import DecodePlan, decode_masks
B, S = 64, 512
ls = [torch.from_numpy(np.tile(synth.logits(8, S, S, seed=s), (B // 8, 1, 1))).cuda() for s in range(4)]
boxes = np.tile(np.array([0, 0, S, S], np.int32), (B, 1))
plan = DecodePlan(boxes, "cuda"); out = torch.empty(B * S * S, dtype=torch.uint8, device="cuda")
for i in range(5): decode_masks(ls[i % 4], out=out, plan=plan)
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
big = torch.randn(8192, 8192, device="cuda")
for _ in range(4): big @ big          # ~tens of ms of queued work: the decode launches below are not host-bound
e0.record()
for i in range(50): decode_masks(ls[i % 4], out=out, plan=plan)
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 50 * 1e3
print(f"slots {os.environ.get('ISG_DEC_SLOTS','default')}: decode {us:.1f} us  {B*S*S*5/us/1e3:.0f} GB/s")
# practical ceiling at this size: a plain device copy moving the same 84 MB (42 MB read + 42 MB write)
srcs = [torch.empty(B * S * S * 5 // 8, dtype=torch.float32, device="cuda") for _ in range(4)]
dsts = [torch.empty_like(s) for s in srcs]
for i in range(5): dsts[i % 4].copy_(srcs[i % 4])
for _ in range(4): big @ big
e0.record()
for i in range(50): dsts[i % 4].copy_(srcs[i % 4])
e1.record(); torch.cuda.synchronize()
us = e0.elapsed_time(e1) / 50 * 1e3
print(f"torch copy of the same {B*S*S*5/1e6:.0f} MB: {us:.1f} us  {B*S*S*5/us/1e3:.0f} GB/s")
