import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")
REF_PATH = os.environ.get("REF_PATH", "/root/reference")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden_state():
    """Calibrated reference state_dict written by tests/golden/make_golden.py."""
    z = np.load(os.path.join(GOLDEN, "segment3_state.npz"))
    return {k: torch.from_numpy(z[k]) for k in z.files}


@pytest.fixture(scope="session")
def golden_logits():
    z = np.load(os.path.join(GOLDEN, "segment3_golden.npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def reference_module():
    """The unmodified reference model/segment.py, when /root/reference is present."""
    path = os.path.join(REF_PATH, "model", "segment.py")
    if not os.path.exists(path):
        pytest.skip("reference tree not present on this machine")
    import importlib.util
    spec = importlib.util.spec_from_file_location("ref_segment_for_tests", path)
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod


def randomized_state(in_channel, seed):
    """A well-conditioned random state for any in_channel: our Segment's seeded init, every
    bias/BN/PReLU path perturbed, BN statistics calibrated with train-mode passes (torch path)."""
    from instancesegmentation_b200 import Segment, synth
    torch.manual_seed(seed)
    m = Segment(in_channel)
    g = torch.Generator().manual_seed(seed + 1)
    with torch.no_grad():
        for name, p in m.named_parameters():
            if name.endswith("act.weight") or name.endswith("prelu.weight"):
                p.copy_(torch.empty_like(p).uniform_(0.05, 0.45, generator=g))
            elif p.dim() == 1 and name.endswith(".weight"):
                p.copy_(torch.empty_like(p).uniform_(0.8, 1.2, generator=g))
            elif name.endswith(".bias"):
                p.copy_(torch.empty_like(p).normal_(0, 0.1, generator=g))
    for mod in m.modules():
        if isinstance(mod, torch.nn.BatchNorm2d):
            mod.momentum = None
            mod.reset_running_stats()
    m.train()
    with torch.no_grad():
        for i in range(4):
            m(torch.from_numpy(synth.batch(4, 128, 128, seed=100 + 4 * i, channels=in_channel)))
    m.eval()
    return m
