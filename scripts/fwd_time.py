"""Forward time of batch B at SxS (bf16 mode): mean over 20 forwards after warm-up, CUDA events."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
B = int(os.environ.get("B", 64)); S = int(os.environ.get("S", 512))
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = os.environ.get("PREC", "bf16"), B
m = m.cuda().eval()
xs = [torch.from_numpy(np.tile(synth.batch(4, S, S, seed=s), (B // 4, 1, 1, 1))).cuda() for s in (1, 2, 3)]
with torch.no_grad():
    for i in range(5): m(xs[i % 3])
    torch.cuda.synchronize()
    best = 1e9
    for rep in range(3):
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(20): m(xs[i % 3])
        e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1) / 20)
print(f"forward B={B} S={S}: {best*1000:.1f} us  ({B/best*1000:.0f} img/s)")
