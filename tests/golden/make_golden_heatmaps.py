"""Golden vectors for the keypoint heat maps from the UNMODIFIED reference function.

``train_instance.py`` cannot be imported (ymlib / imgaug are absent), but ``keypoint2heatmaps``
(train_instance.py:33-68) only needs numpy, math, ``ORDER_PART_NAMES`` (:25-30) and ymlib's
``key_combine``.  This script parses the reference file with ``ast``, executes exactly those two
definitions from it and supplies ``key_combine = lambda name, kind: name`` -- the function uses
``key_combine`` only to build dictionary keys, and the same stub builds the input dictionaries here,
so any injective stand-in gives the function's own output.

Run in the dev container only (needs /root/reference):
    python tests/golden/make_golden_heatmaps.py
Writes tests/golden/heatmaps_golden.npz: per case ``kp`` [17,3] float32 (x, y, visible),
``shape`` (h, w), ``sigma``, ``threshold`` and the reference's 17 heat maps [17,h,w] float32.
"""
import ast
import math
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
REF = os.environ.get("REF_PATH", "/root/reference")


def load_reference_function():
    src = open(os.path.join(REF, "train_instance.py")).read()
    tree = ast.parse(src)
    keep = [n for n in tree.body
            if (isinstance(n, ast.FunctionDef) and n.name == "keypoint2heatmaps")
            or (isinstance(n, ast.Assign) and getattr(n.targets[0], "id", "") == "ORDER_PART_NAMES")]
    assert len(keep) == 2, "reference layout changed"
    ns = {"np": np, "math": math, "key_combine": lambda name, kind: name}
    exec(compile(ast.Module(body=keep, type_ignores=[]), "train_instance.py", "exec"), ns)
    return ns["keypoint2heatmaps"], ns["ORDER_PART_NAMES"]


def to_dict(kp, names):
    """[17,3] (x, y, flag) -> the annotation dictionary the reference reads; flag 0 = key missing,
    1 = visible, 2 = present but not visible"""
    d = {}
    for (x, y, f), name in zip(kp, names):
        if f == 0:
            continue
        d[name] = {"status": "vis" if f == 1 else "hide", "point": (float(x), float(y))}
    return d


def cases():
    rng = np.random.default_rng(2024)
    out = []
    for h, w, sigma, thr in ((120, 160, 10, 0.01), (64, 48, 10, 0.01), (480, 480, 10, 0.01), (96, 96, 4, 0.05)):
        kp = np.zeros((17, 3), np.float32)
        kp[:, 0] = rng.uniform(-30, w + 30, 17)
        kp[:, 1] = rng.uniform(-30, h + 30, 17)
        kp[:, 2] = rng.choice([0, 1, 1, 1, 2], 17)
        kp[0] = (0.0, 0.0, 1)                    # corner, integer coordinates
        kp[1] = (w - 1, h - 1, 1)                # the last row/column is never written (arange excludes x_max)
        kp[2] = (w / 2 + 0.5, h / 2 + 0.25, 1)
        kp[3] = (-25.0, h / 3, 1)                # centre outside, disk reaches in
        kp[4] = (w + 40.0, 5.0, 1)               # fully outside
        out.append((kp.astype(np.float32), (h, w), sigma, thr))
    return out


def main():
    fn, names = load_reference_function()
    blob = {}
    for i, (kp, shape, sigma, thr) in enumerate(cases()):
        maps = np.stack(fn(to_dict(kp, names), shape, sigma=sigma, threshold=thr)).astype(np.float32)
        blob[f"kp{i}"] = kp
        blob[f"meta{i}"] = np.array([shape[0], shape[1], sigma, thr], np.float64)
        blob[f"maps{i}"] = maps
        print(i, shape, "nonzero", int((maps != 0).sum()), "max", float(maps.max()))
    blob["n"] = np.array(len(cases()))
    np.savez_compressed(os.path.join(HERE, "heatmaps_golden.npz"), **blob)


if __name__ == "__main__":
    main()
