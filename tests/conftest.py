"""pytest configuration: registers the `gpu` marker and puts the drop-in package
(`realtime-depth-estimation-nconv_b200/`, which provides `models.*` and `nconv_b200.*`) and the
repo root (for `oracle.*`) on sys.path."""
import os
import sys

import numpy as np
import pytest
import torch

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PKG = os.path.join(ROOT, "realtime-depth-estimation-nconv_b200")
for p in (ROOT, PKG):
    if p not in sys.path:
        sys.path.insert(0, p)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# fp32 parity bound of BASELINE.json's north_star: max-rel 1e-4, with |ref| floored at 1e-3 so
# that exact-zero confidence regions stay meaningful.
REL_TOL = 1e-4


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def pytest_collection_modifyitems(config, items):
    if torch.cuda.is_available():
        return
    skip = pytest.mark.skip(reason="no CUDA device")
    for item in items:
        if "gpu" in item.keywords:
            item.add_marker(skip)


@pytest.fixture(scope="session")
def golden():
    def load(name):
        return np.load(os.path.join(GOLDEN, name), allow_pickle=False)
    return load


def T(a):
    return torch.from_numpy(np.ascontiguousarray(a))


def report(test: str, **values) -> None:
    """Append achieved parity figures (max-rel errors, noise ratios ...) to gpurun_out/parity_report.jsonl so a
    regression inside the tolerance is still visible; the file travels back from the GPU box and a copy of each
    round's run is committed under profiles/."""
    import json
    d = os.path.join(ROOT, "gpurun_out")
    try:
        os.makedirs(d, exist_ok=True)
        with open(os.path.join(d, "parity_report.jsonl"), "a") as f:
            f.write(json.dumps({"test": test, **{k: (float(v) if isinstance(v, (int, float)) else v) for k, v in values.items()}}) + "\n")
    except OSError:
        pass
    print("[parity] %s %s" % (test, values))
