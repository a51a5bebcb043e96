"""Generate the committed fixtures in tests/golden/ from the UNMODIFIED reference.

Run in the dev container only (needs /root/reference):
    python tests/golden/make_golden.py

Writes
  segment3_state.npz   calibrated state_dict of reference ``Segment(3)`` (seed 0)
  segment3_golden.npz  reference logits (+ per-tap checksums) on seeded synthetic inputs
The state follows SURVEY.md section 8(c): reference ``weights_init`` under
``torch.manual_seed(0)``; conv/BN biases, BN gamma/beta and PReLU slopes perturbed so every
bias / mean / slope path is exercised; BN running statistics calibrated by 8 train-mode
passes (cumulative average) on the synthetic generator; then ``eval()``.
"""
import hashlib
import importlib.util
import json
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from instancesegmentation_b200 import synth  # noqa: E402

REF = os.environ.get("REF_PATH", "/root/reference")


def load_reference():
    spec = importlib.util.spec_from_file_location("ref_segment", os.path.join(REF, "model/segment.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def calibrated_reference(in_channel=3, seed=0):
    ref = load_reference()
    torch.manual_seed(seed)
    m = ref.Segment(in_channel)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("act.weight") or name.endswith("prelu.weight"):
                p.copy_(torch.empty_like(p).uniform_(0.05, 0.45, generator=g))
            elif name.endswith("bn.weight") or name.endswith("convs.2.weight") and p.dim() == 1:
                p.copy_(torch.empty_like(p).uniform_(0.8, 1.2, generator=g))
            elif name.endswith(".bias"):
                p.copy_(torch.empty_like(p).normal_(0, 0.1, generator=g))
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.momentum = None
            mod.reset_running_stats()
    m.train()
    with torch.no_grad():
        for i in range(8):
            m(torch.from_numpy(synth.batch(4, 256, 256, seed=100 + 4 * i, channels=in_channel)))
    m.eval()
    return m


CASES = {  # name -> (B, H, W, seed, kind)
    "a_1x64x64": (1, 64, 64, 1234, "natural"),
    "b_2x96x160": (2, 96, 160, 1300, "natural"),
    "c_1x128x128_noise": (1, 128, 128, 1400, "noise"),
}


def main():
    m = calibrated_reference()
    sd = m.state_dict()
    np.savez(os.path.join(HERE, "segment3_state.npz"), **{k: v.numpy() for k, v in sd.items()})
    out, meta = {}, {"keys_sha256": hashlib.sha256("\n".join(sd.keys()).encode()).hexdigest(), "cases": {}}
    taps = {}
    hooks = [getattr(m, n).register_forward_hook(lambda mod, i, o, n=n: taps.__setitem__(n, o))
             for n in ("init_conv", "bottle1_x", "bottle2_x", "bottle3_x", "bottle4_3", "bottle5_2", "bottle6_1")]
    for name, (B, H, W, seed, kind) in CASES.items():
        x = torch.from_numpy(synth.batch(B, H, W, seed=seed, kind=kind))
        with torch.no_grad():
            y = m(x)
        out[name] = y.numpy()
        meta["cases"][name] = dict(B=B, H=H, W=W, seed=seed, kind=kind,
                                   logit_std=float(y.std()), logit_absmax=float(y.abs().max()),
                                   taps={k: [float(v.double().sum()), float(v.double().abs().sum())]
                                         for k, v in taps.items()})
    for h in hooks:
        h.remove()
    np.savez(os.path.join(HERE, "segment3_golden.npz"), **out)
    with open(os.path.join(HERE, "segment3_golden.json"), "w") as f:
        json.dump(meta, f, indent=1)
    print(json.dumps({k: (v["logit_std"], v["logit_absmax"]) for k, v in meta["cases"].items()}))


if __name__ == "__main__":
    main()
