"""CPU suite: the kernels' token-conversion code compiled for the host, the generators, and the
C ABI library's loadability / exported symbols (no compute without a GPU)."""
import ctypes as C
import os
import random
import re

import numpy as np
import pytest

from conftest import ROOT


@pytest.fixture(scope="module")
def emul():
    from dmlc_core_b200 import build
    lib = C.CDLL(build.build_emul())
    lib.emul_strtof_bits.restype = C.c_uint32
    lib.emul_strtof_bits.argtypes = [C.c_char_p, C.POINTER(C.c_int64)]
    lib.emul_strtouint.restype = C.c_uint64
    lib.emul_strtouint.argtypes = [C.c_char_p, C.c_int, C.POINTER(C.c_int)]
    lib.emul_atoll.restype = C.c_int64
    lib.emul_atoll.argtypes = [C.c_char_p]
    return lib


def test_device_token_code_matches_oracle(emul, oracle):
    rnd = random.Random(21)
    toks = ["0." + "0" * 64, "-0." + "0" * 70 + "e2", "1e39", "56018346692", "-", "e", ""]
    for _ in range(100000):
        toks.append("".join(rnd.choice("0123456789" * 3 + "+-.eE") for _ in range(rnd.randint(1, 26))))
    for t in toks:
        n = C.c_int64()
        got = (emul.emul_strtof_bits(t.encode(), C.byref(n)), n.value)
        assert got == oracle.strtof_bits(t), t
    for _ in range(50000):
        t = "".join(rnd.choice("0123456789" * 4 + "+-. ") for _ in range(rnd.randint(0, 24)))
        for w in (4, 8):
            th = C.c_int()
            v = emul.emul_strtouint(t.encode(), w, C.byref(th))
            ov, othrew = oracle.strtouint(t, w)
            assert bool(th.value) == othrew and (othrew or v == ov), (t, w)
        assert emul.emul_atoll(t.encode()) == oracle.atoll(t), t


def test_window_fast_paths_match_oracle(emul, oracle):
    """ToFloatFast / ToUintFast (register-window fast paths with fallback) against the oracle, with
    the token at every alignment, followed by each kind of terminator, near the end of the chunk."""
    emul.emul_fast_strtof_bits.restype = C.c_uint32
    emul.emul_fast_strtof_bits.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.POINTER(C.c_int)]
    emul.emul_fast_strtouint.restype = C.c_uint64
    emul.emul_fast_strtouint.argtypes = [C.c_char_p, C.c_uint32, C.c_uint32, C.c_int,
                                         C.POINTER(C.c_int), C.POINTER(C.c_int)]
    rnd = random.Random(77)
    terms = [",", " ", "\n", ":", "\t", "\r\n", "x", ",,", " 5", ":7", ""]
    fast_f = fast_u = total = 0
    toks = ["-7.60915", "0.137604", "7", "-", "+", ".", "-.5", "5.", "1234567.1234567", "12345678.5",
            "0.12345678", "16777217", "1e5", "2.5e-3", "9999999.9999999", "-0", "007.500", " 5", "\t1",
            "0.0000001", "1.", "+1.5", "--5", "1.2.3", "0.", "-1234567.1234567", "+1234567.1234567",
            # exactly eight fraction digits (what %.6g prints for 0.00123456), nine, and long forms
            "0.00123456", "-0.00123456", "+0.00999999", "12.12345678", "123456.12345678",
            "1234567.12345678", "0.123456789", "0.99999999", "0.999999999", "9.00000001", ".12345678"]
    for _ in range(4000):
        toks.append("%s%d.%s" % (rnd.choice(["", "-", "+"]), rnd.randint(0, 10 ** rnd.randint(0, 7)),
                                 "".join(rnd.choice("0123456789") for _ in range(rnd.choice([7, 8, 8, 8, 9])))))
    for _ in range(60000):
        k = rnd.random()
        if k < 0.4:
            toks.append("%.6g" % rnd.uniform(-10, 10))
        elif k < 0.6:
            toks.append("%.7g" % rnd.uniform(0, 1e6))
        elif k < 0.8:
            toks.append("".join(rnd.choice("0123456789") for _ in range(rnd.randint(1, 9))) +
                        rnd.choice(["", ".", "." + "".join(rnd.choice("0123456789") for _ in range(rnd.randint(0, 9)))]))
        else:
            toks.append("".join(rnd.choice("0123456789" * 3 + "+-.eE") for _ in range(rnd.randint(1, 20))))
    # long integer parts (ids, timestamps): eight to sixteen sign + digit bytes, with and without fraction
    toks += ["16777216", "16777217", "99999999", "123456789", "1609459200", "-1609459200", "+999999999999999",
             "1234567890123456", "1609459200.5", "123456789.123456", "00000000123", "-0000000012.5"]
    for _ in range(6000):
        toks.append(rnd.choice(["", "", "-", "+"]) +
                    "".join(rnd.choice("0123456789") for _ in range(rnd.randint(7, 16))) +
                    rnd.choice(["", "", ".", "." + "".join(rnd.choice("0123456789") for _ in range(rnd.randint(1, 8)))]))
    for tok in toks:
        lead = "," * rnd.randint(0, 9)
        tail = rnd.choice(terms) + ("123.5,\n" * rnd.randint(0, 4))
        buf = (lead + tok + tail).encode()
        took = C.c_int()
        got = emul.emul_fast_strtof_bits(buf, len(buf), len(lead), C.byref(took))
        want, _ = oracle.strtof_bits((tok + tail).encode())
        assert got == want, (tok, tail, lead)
        fast_f += took.value
        total += 1
    # on the bench's token shape (%.6g, other text following) the window path is THE path
    fast_typ = 0
    for i in range(5000):
        tok = "%.6g" % rnd.uniform(-10, 10)
        buf = ("," * (i % 16) + tok + ",-3.25,7.5,0.001,1,2,3\n").encode()
        took = C.c_int()
        emul.emul_fast_strtof_bits(buf, len(buf), i % 16, C.byref(took))
        fast_typ += took.value
    assert fast_typ >= 4950, fast_typ
    n_u = 0
    for _ in range(60000):
        tok = "".join(rnd.choice("0123456789") for _ in range(rnd.randint(0, 27)))
        if rnd.random() < 0.1:
            tok = rnd.choice(["+", "-", " ", "."]) + tok
        lead = " " * rnd.randint(0, 9)
        tail = rnd.choice([":", " ", "\n", ":0.5 ", ".5", "e", ""]) + ("9:1 " * rnd.randint(0, 6))
        buf = (lead + tok + tail).encode()
        for w in (4, 8):
            th, took = C.c_int(), C.c_int()
            v = emul.emul_fast_strtouint(buf, len(buf), len(lead), w, C.byref(th), C.byref(took))
            ov, othrew = oracle.strtouint((tok + tail).encode(), w)
            assert bool(th.value) == othrew and (othrew or v == ov), (tok, tail, w)
            fast_u += took.value
            n_u += 1
    assert fast_u > n_u // 4


def test_digitchar_class(emul):
    want = set(b"0123456789+-.eE")
    for c in range(256):
        assert bool(emul.emul_is_digitchar(c)) == (c in want), c


def test_generators_are_deterministic_and_wellformed(oracle):
    from dmlc_core_b200 import datagen
    for kind, fmt, w, lc in (("libsvm", "libsvm", 4, -1), ("csv", "csv", 4, 0), ("libfm", "libfm", 8, -1),
                             ("libsvm_qid", "libsvm", 4, -1), ("criteo", "libsvm", 8, -1)):
        a, rows = datagen.generate(kind, 7, 300 << 10)
        b, _ = datagen.generate(kind, 7, 300 << 10)
        assert a.tobytes() == b.tobytes()
        assert a[-1] == 10
        res = oracle.parse_buffer(fmt, a.tobytes(), w, lc)
        assert res["error"] == 0 and len(res["label"]) == rows
        if kind == "libsvm_qid":
            assert len(res["qid"]) == rows and len(res["weight"]) == rows and len(res["value"]) == 0
            assert np.all(np.diff(res["qid"].astype(np.int64)) >= 0)
        if kind == "csv":
            assert len(res["index"]) == rows * 128


def test_c_abi_library_exports_header_symbols():
    """libdmlc_b200.so loads without a GPU and exports everything include/*.h declares."""
    from dmlc_core_b200 import build, capi
    build.build_cuda()
    lib = capi.load()
    with open(os.path.join(ROOT, "include", "dmlc_b200_parser.h")) as f:
        header = f.read()
    declared = set(re.findall(r"\b(b200p_[a-z_0-9]+)\s*\(", header))
    assert declared, "no declarations found"
    assert declared == set(capi.EXPORTS)
    for name in declared:
        assert hasattr(lib, name), name
    assert b"sm_100a" in lib.b200p_version()


def test_no_gpu_means_loud_failure():
    """Without a CUDA device the product refuses to run (no CPU fallback)."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from dmlc_core_b200 import Parser
    with pytest.raises(RuntimeError):
        Parser("libsvm")


def test_product_never_imports_oracle():
    """The oracle is test infrastructure: nothing under dmlc_core_b200/ may reference it."""
    pkg = os.path.join(ROOT, "dmlc_core_b200")
    for base, _, files in os.walk(pkg):
        for f in files:
            if f == "build.py":
                continue  # building the checker is not using it
            if f.endswith((".py", ".cu", ".cuh", ".h", ".cc", ".c")):
                with open(os.path.join(base, f), errors="replace") as fh:
                    src = fh.read()
                assert "pyoracle" not in src and "liboracle" not in src and "libdmlc_ref" not in src, f
