"""Per-phase CTA timing of every fused launch (ISG_FK_PROF=1), batch 64 at 512x512."""
import os, sys
os.environ["ISG_FK_PROF"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
B = int(sys.argv[1]) if len(sys.argv) > 1 else 64
S = int(sys.argv[2]) if len(sys.argv) > 2 else 512
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = "bf16", B
m = m.cuda().eval()
x = torch.from_numpy(np.tile(synth.batch(4, S, S, seed=1), (B // 4, 1, 1, 1))).cuda()
with torch.no_grad():
    m(x); torch.cuda.synchronize()
    print("---- second pass", file=sys.stderr)
    m(x); torch.cuda.synchronize()
