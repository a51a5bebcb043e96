"""debug: bf16x3 logits with the cluster-resident mega3 launches vs per-block launches (bitwise) and vs the oracle"""
import os, subprocess, sys
import numpy as np
code = r'''
import sys, numpy as np, torch
sys.path.insert(0, ".")
from instancesegmentation_b200 import Segment, synth
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = "bf16x3", 3
m = m.cuda().eval()
x = torch.from_numpy(synth.batch(3, 512, 512, seed=3)).cuda()
with torch.no_grad(): y = m(x)
torch.cuda.synchronize(); np.save(sys.argv[1], y.cpu().numpy())
'''
outs = []
for i, env in enumerate(({}, {"ISG_FK_NO_MEGA": "1"})):
    subprocess.run([sys.executable, "-c", code, f"/tmp/y{i}.npy"], check=True, env=dict(os.environ, ISG_FK_NO_TUNE="1", **env))
    outs.append(np.load(f"/tmp/y{i}.npy"))
print("mega3 vs per-block launches: equal", np.array_equal(outs[0], outs[1]), "max diff", float(np.abs(outs[0] - outs[1]).max()))
