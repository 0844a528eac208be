import json
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


@pytest.fixture(scope="session")
def golden():
    with open(os.path.join(GOLDEN, "golden.json")) as f:
        return json.load(f)


@pytest.fixture(scope="session")
def images():
    with np.load(os.path.join(GOLDEN, "images.npz")) as z:
        return {k: z[k] for k in z.files}


@pytest.fixture(scope="session")
def golden_lists():
    with np.load(os.path.join(GOLDEN, "keypoints.npz")) as z:
        return {k: z[k] for k in z.files}


def image_for(images, stem, mode):
    """mode 'cvt' = imread COLOR + cvtColor (DoG, Pyramid_Test); 'gray' = IMREAD_GRAYSCALE (Harris, RotateImgTest)."""
    return images[stem] if stem in images else images[f"{stem}_{mode}"]


@pytest.fixture(scope="session")
def ctx():
    import visualslam_b200 as vs
    c = vs.Context(0)
    yield c
    c.close()
