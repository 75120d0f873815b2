import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name), allow_pickle=False)


@pytest.fixture(scope="session")
def golden_c1():
    return load_golden("c1_640x480.npz")


@pytest.fixture(scope="session")
def golden_adv():
    return load_golden("adversarial.npz")


@pytest.fixture(scope="session")
def golden_kat():
    return load_golden("matcher_kat.npz")


def assert_kp_equal(kp, ref, float_rtol=0.0):
    """Keypoint records equal: positions/octave exact; response/angle bit-equal by default
    (north_star allows 1e-4 relative on Harris/orientation floats — pass float_rtol for that)."""
    assert len(kp) == len(ref), f"count {len(kp)} != {len(ref)}"
    assert np.array_equal(kp["octave"], ref["octave"])
    assert np.array_equal(kp["x"], ref["x"]) and np.array_equal(kp["y"], ref["y"])
    if float_rtol == 0.0:
        assert np.array_equal(kp["response"], ref["response"])
        assert np.array_equal(kp["angle"], ref["angle"])
    else:
        np.testing.assert_allclose(kp["response"], ref["response"], rtol=float_rtol)
        np.testing.assert_allclose(kp["angle"], ref["angle"], rtol=float_rtol)


def matches_as_rows(m):
    return np.stack([m["query_idx"], m["train_idx"], m["distance"]], 1).astype(np.int32).reshape(-1, 3)
