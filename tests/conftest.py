import os
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "miotts-llama.cpp_b200"))

GOLDEN = os.path.join(ROOT, "tests", "golden")


def pytest_configure(config):
    config.addinivalue_line("markers", "gpu: needs a B200 (run with -m gpu on the GPU box)")


def snr_db(test, ref):
    test = np.asarray(test, dtype=np.float64)
    ref = np.asarray(ref, dtype=np.float64)
    err = np.sum((test - ref) ** 2)
    return float(10.0 * np.log10(np.sum(ref ** 2) / max(err, 1e-300)))


@pytest.fixture(scope="session")
def assets():
    from miocodec_b200 import synth
    return synth.ensure_assets()


@pytest.fixture(scope="session")
def weights():
    from miocodec_b200 import synth
    hp, tensors = synth.synth_tensors(0)
    return hp, dict(tensors)


def load_golden(name):
    return np.load(os.path.join(GOLDEN, name + ".npz"))


@pytest.fixture(scope="session")
def codec(assets):
    """One GPU context shared by the gpu tests (tcgen05 path)."""
    from miocodec_b200 import binding
    c = binding.MioCodec(assets[0], 0)
    yield c
    c.close()
