"""Where the library-backed training step (bench.py --workload cfg5) spends its GPU time: top kernels of one
eager step of 16 images at 512x512 under torch.profiler.  Usage: python scripts/train_profile.py [channels_last]"""
import sys

import torch
from torch.profiler import ProfilerActivity, profile

sys.path.insert(0, ".")
from instancesegmentation_b200 import Segment, synth  # noqa: E402
from instancesegmentation_b200.train_step import FlatBucket, bce_sigmoid_loss  # noqa: E402

cl = len(sys.argv) > 1 and sys.argv[1] == "channels_last"
if len(sys.argv) > 1 and sys.argv[1] == "benchmark":
    torch.backends.cudnn.benchmark = True
dev = torch.device("cuda")
torch.manual_seed(0)
model = Segment(3).to(dev).train()
bucket = FlatBucket(model, grads_as_views=True)
x = torch.from_numpy(synth.batch(16, 512, 512, seed=1)).to(dev)
y = (torch.from_numpy(synth.logits(16, 512, 512, seed=2)) > 0.8).float()[:, None].to(dev)
if cl:
    x = x.contiguous(memory_format=torch.channels_last)
    for prm in model.parameters():
        if prm.dim() == 4:
            prm.data = prm.data.contiguous(memory_format=torch.channels_last)
    bucket = None


def step():
    if bucket is not None:
        bucket.zero_grads()
    else:
        model.zero_grad(set_to_none=True)
    with torch.autocast("cuda", dtype=torch.bfloat16):
        out = model(x)
    loss = bce_sigmoid_loss(out, y)
    loss.backward()


for _ in range(6):
    step()
torch.cuda.synchronize()
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
e0.record()
for _ in range(5):
    step()
e1.record()
torch.cuda.synchronize()
print(f"step (forward+loss+backward): {e0.elapsed_time(e1) / 5:.2f} ms  channels_last={cl}")
with profile(activities=[ProfilerActivity.CUDA, ProfilerActivity.CPU]) as prof:
    step()
    torch.cuda.synchronize()
print(prof.key_averages().table(sort_by="self_cuda_time_total", row_limit=28, max_name_column_width=70))
