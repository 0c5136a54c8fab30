"""The bounds-checked build (-DB200P_CHECKED) under the parity tests.  compute-sanitizer is not
available on the GPU pool, so out-of-bounds queue writes / stores and any disagreement between
what pass 1 counted and what pass 2 produced are caught by the kernels' own checks instead: the
same parity tests run in a subprocess against libdmlc_b200_checked.so, where a violation turns
into an error return (and so into a test failure) rather than silent corruption."""
import os
import subprocess
import sys

import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def test_parity_subset_on_the_checked_build():
    from dmlc_core_b200 import build
    lib = build.build_cuda_checked()
    env = dict(os.environ)
    env["DMLC_B200_LIB"] = lib
    out = subprocess.run(
        [sys.executable, "-m", "pytest", os.path.join(ROOT, "tests", "test_gpu_parity.py"), "-x", "-q",
         "-m", "gpu", "-k", "golden or fuzz or corpus_device or edge or streaming or max_index",
         "-p", "no:cacheprovider"],
        cwd=ROOT, env=env, stdout=subprocess.PIPE, stderr=subprocess.STDOUT, text=True, timeout=1500)
    tail = out.stdout[-3000:]
    assert out.returncode == 0, tail
    assert " passed" in tail


def test_checked_build_identifies_itself():
    import ctypes
    from dmlc_core_b200 import build
    lib = ctypes.CDLL(build.build_cuda_checked())
    lib.b200p_version.restype = ctypes.c_char_p
    assert b"checked" in lib.b200p_version()
