import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")

# tolerance of BASELINE.json's north star: state and covariance within 1e-9 relative
# (max-norm and covariance-scaled)
TOL = 1e-9


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")
    config.addinivalue_line("markers", "reference: needs /root/reference (this container only)")


def golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def surrogate():
    """events.csv lines + detections sidecar + the reference replay's per-step outputs."""
    import gzip
    from oracle.ekf_oracle import Detection
    d = os.path.join(GOLDEN, "surrogate")
    with gzip.open(os.path.join(d, "events.csv.gz"), "rt") as fh:
        lines = fh.readlines()
    z = np.load(os.path.join(d, "detections.npz"))
    dets = {}
    for i, nm in enumerate(z["names"]):
        lo, hi = z["offsets"][i], z["offsets"][i + 1]
        dets[str(nm)] = [Detection(int(z["tag_id"][k]), z["pose_t"][k].reshape(3, 1), float(z["pose_err"][k]))
                         for k in range(lo, hi)]
    ref = np.load(os.path.join(d, "replay_ref.npz"))
    return lines, dets, ref


def assert_parity(mu, p, mu_ref, p_ref, tol=TOL, what=""):
    from oracle.ekf_oracle import parity_errors
    e = parity_errors(mu, p, mu_ref, p_ref)
    assert max(e) <= tol, f"{what}: mu {e[0]:.3e}  P max-norm {e[1]:.3e}  P cov-scaled {e[2]:.3e} > {tol}"
    return e
