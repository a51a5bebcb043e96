"""debug: per-tap and logits error of a precision mode against the oracle (GPU box)"""
import sys, os
import numpy as np, torch
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from instancesegmentation_b200 import Segment, synth
from oracle.segment_oracle import segment_forward

prec = sys.argv[1] if len(sys.argv) > 1 else "bf16x3"
B, H, W = (int(v) for v in (sys.argv[2] if len(sys.argv) > 2 else "2x96x160").split("x"))
z = np.load("tests/golden/segment3_state.npz")
sd = {k: torch.from_numpy(z[k]) for k in z.files}
m = Segment(3); m.load_state_dict(sd); m.precision, m.max_batch = prec, B
m = m.cuda().eval()
x = torch.from_numpy(synth.batch(B, H, W, seed=4321))
taps = {}
want = segment_forward(sd, x, taps)
with torch.no_grad():
    got = m(x.cuda()).cpu()
mine = next(iter(m._engines.values())).taps(B)
for k, t in mine.items():
    r = float((t - taps[k]).abs().max() / taps[k].abs().max().clamp_min(1e-6))
    print(f"  tap {k:14s} rel-max {r:.2e}")
d = (got - want).abs()
print(f"{prec} {B}x{H}x{W}: logits max {float(d.max()):.3e} mean {float(d.mean()):.3e}")
