import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run with -m gpu on the B200 box)")


@pytest.fixture(scope="session")
def golden_small():
    import numpy as np

    return np.load(os.path.join(ROOT, "tests", "golden", "golden_small.npz"))


@pytest.fixture(scope="session")
def golden_digests():
    import json

    with open(os.path.join(ROOT, "tests", "golden", "golden_digests.json")) as fh:
        return json.load(fh)


@pytest.fixture(scope="session")
def gb():
    """The package with its C-ABI library built in-tree."""
    import libgaussianblur_b200 as g
    from libgaussianblur_b200 import build

    build.build()
    g.lib()
    return g


@pytest.fixture(scope="session")
def refcheck():
    """Checker used by the parity tests: the compiled reference when oracle/_ref travelled with
    the repo, else the pinned numpy restatement (tolerance is 1 LSB either way)."""
    from oracle import gblur_oracle as go
    from oracle import ref_lib

    class Check:
        have_ref = ref_lib.available()

        @staticmethod
        def blur(img, sigma, alpha):
            if img.ndim == 2:
                return ref_lib.blur_plane_u8(img, sigma) if ref_lib.available() else go.gaussianblur(img, sigma, alpha)
            return ref_lib.gaussianblur(img, sigma, alpha) if ref_lib.available() else go.gaussianblur(img, sigma, alpha)

        @staticmethod
        def rowpass(img, ch, sigma):
            if ref_lib.available():
                return ref_lib.rowpass_plane(img, ch, sigma)
            return go.blur_plane(img[..., ch].astype("float64"), sigma, return_intermediate=True)[1]

    return Check
