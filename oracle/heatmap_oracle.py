"""TEST INFRASTRUCTURE ONLY -- numpy restatement of the reference's keypoint heat maps.

Follows ``keypoint2heatmaps`` (reference train_instance.py:33-68) on array input: ``kp`` [17,3]
rows (x, y, flag) in ``ORDER_PART_NAMES`` order (train_instance.py:25-30), flag 1 = key present
with status 'vis' (a map is drawn), anything else = map stays zero.  Arithmetic as the reference
performs it: float64 throughout (integer grids minus Python floats), result stored into a
float32 array.  PINNED: tests/test_oracle.py compares it with tests/golden/heatmaps_golden.npz,
which tests/golden/make_golden_heatmaps.py wrote by executing the reference function itself.
Only tests/ may import this module; the product never does.
"""
import math

import numpy as np

N_PARTS = 17   # len(ORDER_PART_NAMES), train_instance.py:25-30


def keypoint_heatmaps(kp, shape, sigma=10, threshold=0.01):
    kp = np.asarray(kp)
    h, w = shape
    r = math.sqrt(math.log(threshold) * (-sigma ** 2))                   # :35
    out = np.zeros((len(kp), h, w), np.float32)                          # :41
    for i, (x, y, flag) in enumerate(kp):
        if flag != 1:                                                    # :45-47
            continue
        x, y = float(x), float(y)
        x_min, x_max = max(0, int(x - r)), min(w - 1, int(x + r + 1))    # :52-53
        y_min, y_max = max(0, int(y - r)), min(h - 1, int(y + r + 1))    # :54-55
        xs = np.arange(x_min, x_max)                                     # :57 (x_max itself excluded)
        ys = np.arange(y_min, y_max)[:, None]                            # :58
        e = np.exp(-((xs - x) ** 2 + (ys - y) ** 2) / sigma ** 2)        # :60
        idx = np.where(e > threshold)                                    # :62
        out[i, y_min:y_max, x_min:x_max][idx] = e[idx]                   # :64
    return out
