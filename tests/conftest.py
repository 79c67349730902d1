import os
import sys
import tarfile

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)
GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def fd():
    import fdb200
    return fdb200


@pytest.fixture(scope="session")
def shipped_model_dir(tmp_path_factory):
    """The reference's saved PipelineModel, unpacked from the committed archive."""
    d = tmp_path_factory.mktemp("shipped")
    with tarfile.open(os.path.join(GOLDEN, "shipped_model.tar.gz")) as tf:
        tf.extractall(d, filter="data")
    return str(d / "dialogue_classification_model")


@pytest.fixture(scope="session")
def golden():
    import json
    with open(os.path.join(GOLDEN, "golden_vectors.json")) as f:
        return json.load(f)
