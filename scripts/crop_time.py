"""crop_resize at cfg4: 640 windows U(96,640) of 32 source images -> 640 x 1024x1024x3 uint8 (2.0 GB written)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import synth
from instancesegmentation_b200.preprocess import crop_resize
n_img, per, S = 32, 20, 1024
boxes, first = synth.boxes(n_img, per, S, S, seed=7)
src = torch.from_numpy(np.stack([synth.image_u8(S, S, 1234 + i) for i in range(4)])).cuda()[torch.arange(n_img) % 4].contiguous()
win = torch.from_numpy(boxes).cuda(); ii = torch.from_numpy(np.repeat(np.arange(n_img, dtype=np.int32), per)).cuda()
out = torch.empty((n_img * per, S, S, 3), dtype=torch.uint8, device="cuda")
e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
for _ in range(2): crop_resize(src, win, ii, (S, S), out=out)
torch.cuda.synchronize(); e0.record()
for _ in range(5): crop_resize(src, win, ii, (S, S), out=out)
e1.record(); torch.cuda.synchronize()
ms = e0.elapsed_time(e1) / 5
print(f"crop_resize {ms*1e3:.1f} us  {out.numel()/ms/1e6:.0f} GB/s written")
