import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box)")


@pytest.fixture(scope="session")
def golden():
    import json

    with open(os.path.join(GOLDEN, "reference_vectors.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def coarse_box_mesh():
    from blitzdg_b200 import MeshManager

    m = MeshManager()
    m.readMesh(os.path.join(GOLDEN, "coarse_box.msh"))
    return m


@pytest.fixture(scope="session")
def coarse_quads_mesh():
    from blitzdg_b200 import MeshManager

    m = MeshManager()
    m.readMesh(os.path.join(GOLDEN, "coarse_box_quads.msh"))
    return m
