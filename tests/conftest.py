import json
import os
import sys

import numpy as np
import pytest

os.environ["GGPM_ALLOW_TEST_BACKEND"] = "1"      # the CPU tests route the C ABI to tests/_emu_lib.py (ggpm_b200/_lib.py seam)
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN_DIR = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


def golden_names():
    return sorted(f[:-4] for f in os.listdir(GOLDEN_DIR) if f.endswith(".npz"))


def load_golden(name):
    """-> (meta dict, arrays dict, tree_tensors, graph_tensors) with torch.long index tensors."""
    import torch
    z = np.load(os.path.join(GOLDEN_DIR, name + ".npz"))
    meta = json.loads(bytes(z["meta"]).decode())
    def lt(k):
        return torch.from_numpy(z[k].astype(np.int64))
    tree = [lt("tree_fnode"), lt("tree_fmess"), lt("tree_agraph"), lt("tree_bgraph"), lt("tree_cgraph"),
            [tuple(int(x) for x in r) for r in z["tree_scope"]]]
    graph = [lt("graph_fnode"), lt("graph_fmess"), lt("graph_agraph"), lt("graph_bgraph"),
             [tuple(int(x) for x in r) for r in z["graph_scope"]]]
    return meta, z, tree, graph


@pytest.fixture(scope="session")
def root_dir():
    return ROOT


def pytest_sessionfinish(session, exitstatus):
    """measured worst case per tensor of the parity comparisons (tests/_checks.py REPORT) -> gpurun_out/parity_report.json
    (only when a GPU run produced entries; committed under profiles/ per round)"""
    try:
        import _checks
        import torch
        if not _checks.REPORT or not torch.cuda.is_available():
            return
        out_dir = os.path.join(ROOT, "gpurun_out")
        os.makedirs(out_dir, exist_ok=True)
        path = os.path.join(out_dir, "parity_report.json")
        old = {}
        if os.path.exists(path):
            try:
                old = json.load(open(path))
            except Exception:
                old = {}
        old.update(_checks.REPORT)
        with open(path, "w") as f:
            json.dump(old, f, indent=1, sort_keys=True)
    except Exception:
        pass
