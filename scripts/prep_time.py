"""forward_u8 vs forward (fp32) time, batch 64 at 512x512."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment
B, S = 64, 512
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = "bf16", B
m = m.cuda().eval()
rng = np.random.default_rng(0)
us = [torch.from_numpy(rng.integers(0, 256, size=(B, S, S, 3), dtype=np.uint8)).cuda() for _ in range(3)]
xs = [((u.float() / 255 - 0.5) / 0.5).permute(0, 3, 1, 2).contiguous() for u in us]
def t(f, arg):
    with torch.no_grad():
        for i in range(5): f(arg[i % 3])
        torch.cuda.synchronize(); best = 1e9
        for rep in range(3):
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record()
            for i in range(20): f(arg[i % 3])
            e1.record(); torch.cuda.synchronize(); best = min(best, e0.elapsed_time(e1) / 20)
    return best * 1e3
print(f"forward fp32-in {t(m, xs):.1f} us | forward_u8 {t(m.forward_u8, us):.1f} us")
