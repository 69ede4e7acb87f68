import glob
import json
import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def pytest_collection_modifyitems(config, items):
    try:
        import torch
        has_cuda = torch.cuda.is_available()
    except Exception:
        has_cuda = False
    if has_cuda:
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


class Golden:
    def __init__(self, path):
        d = np.load(path)
        self.meta = json.loads(str(d["meta"]))
        self.arrays = {k: d[k] for k in d.files if k != "meta"}
        self.name = self.meta["name"]

    def __getitem__(self, k):
        return self.arrays[k]

    @property
    def weights(self):
        return [self.arrays["w%d" % i] for i in range(len(self.meta["dims"]) - 1)]


def golden_cases(kind):
    out = []
    for f in sorted(glob.glob(os.path.join(GOLDEN_DIR, "*.npz"))):
        g = Golden(f)
        if g.meta["kind"] == kind:
            out.append(g)
    return out


def max_rel_to_peak(a, b):
    """max|a-b| / max|b| -- the metric SURVEY.md 8(c) prescribes for tensors that cross zero."""
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.abs(a - b).max() / max(np.abs(b).max(), 1e-300))


def l2_rel(a, b):
    a = np.asarray(a, dtype=np.float64)
    b = np.asarray(b, dtype=np.float64)
    return float(np.linalg.norm((a - b).ravel()) / max(np.linalg.norm(b.ravel()), 1e-300))
