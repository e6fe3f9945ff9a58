"""Test configuration.  `-m "not gpu"`: oracle vs the reference's golden vectors, host logic, ABI surface (no compute).
`-m gpu`: the parity tests proper — the CUDA path through the C ABI vs the oracle."""
import os
import sys
import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

BUNDLED = ["cube", "gardenlight", "tenryuu", "manfigure", "clumsydragon"]


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (B200); run with -m gpu")


@pytest.fixture(scope="session")
def O():
    """The checkers (oracle/): CPU restatement + (where present) the reference's own headers compiled in place."""
    from oracle import oracle_py
    oracle_py.build(("oracle",))
    return oracle_py


@pytest.fixture(scope="session")
def golden():
    def load(name):
        with np.load(os.path.join(ROOT, "tests", "golden", f"bvh_{name}.npz")) as z:
            return {k: z[k] for k in z.files}
    return load


@pytest.fixture(scope="session")
def ctx():
    """One context (device 0, own stream) for the whole GPU session; fails loudly without the library or a GPU."""
    from openclraytracer_b200 import api
    c = api.Context(0)
    yield c
    c.close()


def masked_nodes(nodes, n):
    """Lanes the reference leaves undefined (inner data[3]) are zeroed before comparing."""
    m = nodes.copy()
    m["data"][n:, 3] = 0
    m["bounds"][n:, 1, 3] = 0
    if n == 1:
        m["data"][:, 0] = 0xFFFFFFFF
    return m


def digest(a):
    import hashlib
    return hashlib.sha256(np.ascontiguousarray(a).view(np.uint8).tobytes()).hexdigest()
