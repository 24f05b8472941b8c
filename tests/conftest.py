import os
import sys

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a CUDA device (run on the B200 box with -m gpu)")


@pytest.fixture(scope="session")
def oracle():
    from oracle import pyoracle
    pyoracle.build()
    return pyoracle


def missing_reference_binary(what):
    """oracle/_ref/* are the reference's own sources compiled in place (make -C oracle ref): git-ignored, but they travel with
    the gpurun snapshot.  Where they MUST exist - the reference tree is here to build them from, or a CUDA device is present
    (the GPU box gets the snapshot with the prebuilt binaries) - their absence is a failure: a lost binary must not turn the
    strongest boundary evidence into a green run with fewer tests.  Elsewhere (a plain checkout without the reference) skip."""
    must = os.path.exists("/root/reference/src/ORBextractor.cc")
    if not must:
        try:
            import torch
            must = torch.cuda.is_available()
        except Exception:
            must = False
    msg = f"{what} not built (make -C oracle ref; needs /root/reference)"
    if must:
        pytest.fail(msg)
    pytest.skip(msg)
