import os, sys
os.environ["ISG_OP_TIMING"] = "1"
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
B = int(os.environ.get("B", 64)); S = int(os.environ.get("S", 512))
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = os.environ.get("PREC", "bf16"), B
m = m.cuda().eval()
x = torch.from_numpy(np.tile(synth.batch(4, S, S, seed=1), (B // 4, 1, 1, 1))).cuda()
with torch.no_grad():
    for i in range(3):
        print("---- pass", i, file=sys.stderr)
        m(x)
