"""Where the e2e step goes: per-step durations of the H2D copy, the compute (forward_u8 + decode) and the D2H copy
inside the three-stream Pipeline, against the step period."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment
from instancesegmentation_b200.pipeline import Pipeline, identity_boxes
B, S, K = 64, 512, 30
depth = int(os.environ.get("DEPTH", 2))
z = np.load("tests/golden/segment3_state.npz")
m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = "bf16", B
m = m.cuda().eval()
rng = np.random.default_rng(0)
hb = [torch.from_numpy(rng.integers(0, 256, size=(B, S, S, 3), dtype=np.uint8)).pin_memory() for _ in range(3)]
mh = [torch.empty(B * S * S, dtype=torch.uint8).pin_memory() for _ in range(depth)]
ah = [torch.empty(B, dtype=torch.int32).pin_memory() for _ in range(depth)]
bh = [torch.empty((B, 4), dtype=torch.int32).pin_memory() for _ in range(depth)]
pipe = Pipeline(m, identity_boxes(B, S, S), depth=depth)
for i in range(5): pipe.submit(hb[i % 3], mh[i % depth], (ah[i % depth], bh[i % depth]))
pipe.finish(); torch.cuda.synchronize()
ev = lambda: torch.cuda.Event(enable_timing=True)
h0, h1, c0, c1, d1 = [[ev() for _ in range(K)] for _ in range(5)]
# instrument by wrapping the streams' work: record around submit's pieces through the slot events
t0, t1 = ev(), ev()
t0.record()
for i in range(K):
    with torch.cuda.stream(pipe.h2d): h0[i].record(pipe.h2d)
    c0[i].record()
    s = pipe.submit(hb[i % 3], mh[i % depth], (ah[i % depth], bh[i % depth]))
    with torch.cuda.stream(pipe.h2d): h1[i].record(pipe.h2d)
    c1[i].record()
    with torch.cuda.stream(pipe.d2h): d1[i].record(pipe.d2h)
pipe.finish(); t1.record(); torch.cuda.synchronize()
per = t0.elapsed_time(t1) / K
h = np.median([h0[i].elapsed_time(h1[i]) for i in range(5, K)])
c = np.median([c0[i].elapsed_time(c1[i]) for i in range(5, K)])
gap = np.median([c1[i - 1].elapsed_time(c1[i]) for i in range(6, K)])
print(f"depth {depth}: step period {per*1e3:.0f} us ({B/per*1e3:.0f} img/s) | H2D stream busy per step {h*1e3:.0f} us (incl. waits) | "
      f"compute stream per step {c*1e3:.0f} us (incl. waits) | compute-end to compute-end {gap*1e3:.0f} us")
