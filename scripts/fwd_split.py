"""Experiment: the 64-image forward as NS independent chains of 64/NS images on NS streams (separate plans),
so that one chain's kernels fill the other's tails.  Compares with the single-chain forward."""
import os, sys, copy
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
B = int(os.environ.get("B", 64)); S = int(os.environ.get("S", 512)); NS = int(os.environ.get("NS", 2))
z = np.load("tests/golden/segment3_state.npz")
def mk(b):
    m = Segment(3); m.load_state_dict({k: torch.from_numpy(z[k]) for k in z.files}); m.precision, m.max_batch = "bf16", b
    return m.cuda().eval()
full = mk(B)
parts = [mk(B // NS) for _ in range(NS)]
streams = [torch.cuda.Stream() for _ in range(NS)]
xs = [torch.from_numpy(np.tile(synth.batch(4, S, S, seed=s), (B // 4, 1, 1, 1))).cuda() for s in (1, 2, 3)]
outs = [torch.empty((B // NS, 1, S, S), device="cuda") for _ in range(NS)]
def run_full(i): full(xs[i % 3])
def run_split(i):
    x = xs[i % 3]
    cur = torch.cuda.current_stream()
    for k in range(NS):
        streams[k].wait_stream(cur)
        with torch.cuda.stream(streams[k]):
            parts[k](x[k * (B // NS):(k + 1) * (B // NS)])
    for k in range(NS): cur.wait_stream(streams[k])
with torch.no_grad():
    for name, fn in (("single chain", run_full), (f"{NS} chains", run_split), ("single chain", run_full), (f"{NS} chains", run_split)):
        for i in range(5): fn(i)
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for i in range(20): fn(i)
        e1.record(); torch.cuda.synchronize()
        t = e0.elapsed_time(e1) / 20
        print(f"{name:>14}: {t*1000:.1f} us per {B} images ({B/t*1000:.0f} img/s)")
    a = full(xs[0]); run_split(0); torch.cuda.synchronize()
