import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def meshes():
    z = np.load(os.path.join(GOLDEN, "meshes.npz"))
    return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden():
    z = np.load(os.path.join(GOLDEN, "golden_xt32.npz"))
    return {k: z[k] for k in z.files}


def xt32_sensor_args(uid="0000"):
    """(elev, az_begin, az_end, H, q_wxyz, t_xyz) of the reference's XT-32 configs, read from the fixture JSON
    with a tiny //-comment stripper (test-side only; the product has its own C++ parser)."""
    import json
    sys.path.insert(0, GOLDEN)
    from make_golden import load_json
    c = load_json(os.path.join(GOLDEN, f"xt32_lidar_{uid}.json"))
    sc = c["device"]["sensorConfig"]
    h = c["channels"]["horizontal"]
    return (np.asarray(c["channels"]["vertical"], np.float32), h["range"]["begin"], h["range"]["end"], h["count"],
            [sc["sensor_to_base"][k] for k in ("qw", "qx", "qy", "qz")],
            [sc["base_to_origin"]["tx"], sc["base_to_origin"]["ty"], sc["sensor_to_base"]["tz"]])
