import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def golden():
    path = os.path.join(ROOT, "tests", "golden", "icp_golden.npz")
    return np.load(path, allow_pickle=False)


def golden_cases(g):
    for name in g["case_names"]:
        name = str(name)
        eps, mi, st, nz = g[f"case/{name}/params"]
        yield dict(
            name=name,
            src=g[str(g[f"case/{name}/src"])], tgt=g[str(g[f"case/{name}/tgt"])],
            init=g[f"case/{name}/init"], epsilon=float(eps), max_iters=int(mi),
            stopping_thresh=float(st), normalize=bool(nz),
            T=g[f"case/{name}/T"], err=float(g[f"case/{name}/err"]),
            corr=g[f"case/{name}/corr"], passes=int(g[f"case/{name}/passes"]),
        )


def reference_module():
    """The reference's modules/icp.py if it is mounted (authoring container only)."""
    ref_root = os.environ.get("LIDAR_SLAM_REF", "/root/reference")
    if not os.path.exists(os.path.join(ref_root, "modules", "icp.py")):
        return None
    import importlib.util
    spec = importlib.util.spec_from_file_location("_ref_icp", os.path.join(ref_root, "modules", "icp.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)
    return mod
