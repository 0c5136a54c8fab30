"""CPU suite: pins the oracle (oracle/dmlc_oracle.c) against the reference's own known-answer
literals, the committed golden RowBlocks, and -- when present -- the compiled reference."""
import json
import os
import random

import numpy as np
import pytest

from conftest import GOLDEN, golden_cases, load_golden
from oracle.pyoracle import same_rowblocks

# reference test/strtonum_test.cc:8-10,18-20,28-30 -- expected value is libc's
REF_TEST_FLOATS = ["1234567901234", "+12345.6789", "-0.00123", "+0123.234e-2",
                   "-234234.123123e20", "3.1029831e+38", "000.123e-28"]
REF_TEST_ULL = ["2147483647", "+12345", "18446744073709551615"]


def test_reference_strtonum_literals(oracle):
    for t in REF_TEST_FLOATS:
        bits, _ = oracle.strtof_bits(t)
        assert bits == int(np.float32(float(t)).view(np.uint32)), t
    for t in REF_TEST_ULL:
        v, threw = oracle.strtouint(t, 8)
        assert not threw and v == int(t)


def test_golden_strtonum(oracle):
    with open(os.path.join(GOLDEN, "strtonum.json")) as f:
        gold = json.load(f)
    for tok, (bits, used) in gold["strtof"].items():
        assert oracle.strtof_bits(tok) == (bits, used), tok
    for tok, (v, threw) in gold["strtouint32"].items():
        assert oracle.strtouint(tok, 4) == (v, bool(threw)), tok
    for tok, (v, threw) in gold["strtouint64"].items():
        assert oracle.strtouint(tok, 8) == (v, bool(threw)), tok


def test_atoll_matches_libc(oracle):
    import ctypes
    libc = ctypes.CDLL(None)
    libc.atoll.restype = ctypes.c_longlong
    libc.atoll.argtypes = [ctypes.c_char_p]
    rnd = random.Random(5)
    toks = ["5", "-5", "+7", "  12", "\t\n 3x", "9223372036854775807", "9223372036854775808",
            "-9223372036854775808", "-9223372036854775809", "99999999999999999999999", "", "-",
            "\v\f\r8", "1 2"]
    for _ in range(2000):
        toks.append("".join(rnd.choice("0123456789" * 3 + "+- \tx") for _ in range(rnd.randint(0, 24))))
    for t in toks:
        assert oracle.atoll(t) == libc.atoll(t.encode()), repr(t)


@pytest.mark.parametrize("case", golden_cases(), ids=lambda c: c["name"])
def test_golden_rowblocks(oracle, case):
    text, exp = load_golden(case)
    got = oracle.parse_buffer(case["format"], text, case["index_bytes"], case["label_column"])
    assert same_rowblocks(got, exp) is None


def _random_text(rnd, fmt):
    def num():
        k = rnd.random()
        if k < 0.3:
            return str(rnd.randint(0, 10 ** rnd.randint(1, 12)))
        if k < 0.6:
            return "%.6g" % rnd.uniform(-100, 100)
        if k < 0.7:
            return "%e" % rnd.uniform(-1e10, 1e10)
        if k < 0.8:
            return rnd.choice(["+1", "-1", ".5", "1.", "-", "e", "1e5e3", "1-2", "+", "+-3"])
        return "".join(rnd.choice("0123456789.eE+-") for _ in range(rnd.randint(1, 8)))
    seps = [" ", " ", " ", "  ", "\t", ":", ":", " : ", ": ", " :", ",", "|", "qid:", " qid:", "x", "::"]
    eols = ["\n", "\n", "\n", "\r\n", "\r", "\n\n"]
    out = []
    for _ in range(rnd.randint(0, 6)):
        line = "".join(num() + rnd.choice(seps) for _ in range(rnd.randint(0, 12)))
        out.append(line + rnd.choice(eols))
    return "".join(out).encode()


@pytest.mark.parametrize("fmt,widths", [("libsvm", (4, 8)), ("libfm", (4, 8)), ("csv", (4,))])
def test_oracle_vs_reference_fuzz(oracle, reference, fmt, widths):
    rnd = random.Random(hash(fmt) & 0xffff)
    for _ in range(4000):
        text = _random_text(rnd, fmt)
        w = rnd.choice(widths)
        lc = rnd.choice([-1, 0, 1, 3])
        a = oracle.parse_buffer(fmt, text, w, lc)
        b = reference.parse_buffer(fmt, text, w, lc)
        assert same_rowblocks(a, b) is None, (fmt, w, lc, text)


def test_oracle_vs_reference_float_fuzz(oracle, reference):
    rnd = random.Random(99)
    for _ in range(50000):
        t = "".join(rnd.choice("0123456789" * 3 + "+-.eE") for _ in range(rnd.randint(1, 26)))
        assert oracle.strtof_bits(t) == reference.strtof_bits(t), t


def test_partition_matches_reference(oracle, reference, tmp_path):
    """InputSplit byte ranges: oracle_partition vs the reference's ResetPartition
    (src/io/input_split_base.cc:30-64) on multi-file inputs with CRLF and blank lines."""
    rnd = random.Random(2)
    paths = []
    for i in range(3):
        lines = []
        for _ in range(rnd.randint(5, 60)):
            lines.append(" ".join("%d:%d" % (rnd.randint(1, 99), rnd.randint(0, 9))
                                  for _ in range(rnd.randint(0, 9))) + rnd.choice(["\n", "\r\n", "\n\n"]))
        p = tmp_path / ("part%d.txt" % i)
        p.write_bytes(("1 " + "1 ".join(lines)).encode())
        paths.append(str(p))
    uri = ";".join(paths)
    total = sum(os.path.getsize(p) for p in paths)
    for nparts in (1, 2, 3, 5, 8, 64):
        joined = b""
        for part in range(nparts):
            mine, b, e = oracle.split_bytes(paths, part, nparts)
            theirs, n, _ = reference.split_bytes(uri, part, nparts, total + 16)
            assert mine == theirs, (nparts, part, b, e, n)
            joined += mine
        assert len(joined) == total


def test_parse_uri_parts_cover_file(oracle, reference, tmp_path):
    """Parser::Create(uri, k, n) over all k returns every row once."""
    from dmlc_core_b200 import datagen
    data, rows = datagen.generate("libsvm", 3, 200 << 10)
    p = tmp_path / "a.svm"
    p.write_bytes(data.tobytes())
    whole = oracle.parse_buffer("libsvm", data.tobytes(), 4)
    assert len(whole["label"]) == rows
    for nparts in (2, 4):
        labels, index = [], []
        for k in range(nparts):
            r = reference.parse_uri(str(p), k, nparts, "libsvm", 4)
            o = oracle.parse_files("libsvm", [str(p)], k, nparts, 4)
            assert same_rowblocks(r, o) is None
            labels.append(r["label"])
            index.append(r["index"])
        assert np.array_equal(np.concatenate(labels), whole["label"])
        assert np.array_equal(np.concatenate(index), whole["index"])
