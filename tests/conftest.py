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
def oracle():
    from oracle.pyoracle import Oracle
    return Oracle()


@pytest.fixture(scope="session")
def reference():
    """The compiled unmodified reference (oracle/_ref); tests that need it skip when absent."""
    from oracle.pyoracle import Reference
    ref = Reference.load()
    if ref is None:
        pytest.skip("oracle/_ref/libdmlc_ref.so not built")
    return ref


def golden_cases():
    with open(os.path.join(GOLDEN, "index.json")) as f:
        return json.load(f)


def load_golden(case):
    name = case["name"]
    with open(os.path.join(GOLDEN, name + ".txt"), "rb") as f:
        text = f.read()
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    exp = {k: z[k] for k in ("offset", "label", "weight", "qid", "field", "index", "value")}
    exp["error"] = int(z["error"])
    return text, exp


@pytest.fixture(scope="session")
def cuda_lib():
    """Builds (if stale) and loads libdmlc_b200.so; GPU tests go through this C ABI only."""
    from dmlc_core_b200 import build, capi
    build.build_cuda()
    return capi.load()
