"""Layer-wise check of the fused bf16 path against the oracle (debug aid, GPU box)."""
import os, sys
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
import numpy as np, torch
from instancesegmentation_b200 import Segment, synth
from oracle.segment_oracle import segment_forward

sizes = [(2, 96, 160), (2, 512, 512), (1, 208, 48), (1, 16, 16)] if len(sys.argv) < 2 else [tuple(map(int, a.split("x"))) for a in sys.argv[1:]]
z = np.load("tests/golden/segment3_state.npz")
sd = {k: torch.from_numpy(z[k]) for k in z.files}
for B, H, W in sizes:
    m = Segment(3); m.load_state_dict(sd, strict=True); m.precision, m.max_batch = "bf16", B
    m = m.cuda().eval()
    x = torch.from_numpy(synth.batch(B, H, W, seed=4321))
    taps = {}
    want = segment_forward(sd, x, taps)
    with torch.no_grad():
        got = m(x.cuda()).cpu()
    torch.cuda.synchronize()
    eng = next(iter(m._engines.values()))
    mine = eng.taps(B)
    print(f"== {B}x{H}x{W}: launches {eng.launches_per_forward}")
    for k, t in mine.items():
        if k not in taps:
            print("  (no oracle tap for", k, ")"); continue
        ref = taps[k]
        if t.shape != ref.shape:
            print("  SHAPE", k, tuple(t.shape), tuple(ref.shape)); continue
        d = (t - ref).abs()
        rel = float(d.max() / ref.abs().max().clamp_min(1e-6))
        flag = "" if rel < 0.05 else "   <<<<<<"
        # where is the worst error?
        idx = np.unravel_index(int(d.argmax()), d.shape)
        print(f"  {k:14s} rel {rel:.2e} mean {float(d.mean()):.2e} refmax {float(ref.abs().max()):.2f} worst@{idx}{flag}")
    d = (got - want).abs()
    idx = np.unravel_index(int(d.argmax()), d.shape)
    print(f"  logits max {float(d.max()):.3e} mean {float(d.mean()):.3e} (ref absmax {float(want.abs().max()):.2f}) worst@{idx} nan={bool(torch.isnan(got).any())}")
    # border vs interior error
    inner = d[..., 1:-1, 1:-1]
    print(f"  interior max {float(inner.max()):.3e}; border-ring max {float(max(d[...,0,:].max(), d[...,-1,:].max(), d[...,:,0].max(), d[...,:,-1].max())):.3e}")
