"""TEST INFRASTRUCTURE: regenerate tests/golden/ from the compiled, unmodified reference.

    python oracle/make_golden.py        (needs oracle/_ref/libdmlc_ref.so, i.e. /root/reference)

For every case it writes <name>.txt (the input bytes) and <name>.npz (the concatenated RowBlock
arrays the reference returned through Parser<IndexType>::Create on that file, plus `error`).
The cases cover the quirk table of SURVEY.md section 8(a) and small slices of the five
BASELINE.json corpus shapes.  tests/golden/strtof.json holds reference strtof bit patterns for
the literals of the reference's test/strtonum_test.cc and SURVEY.md Appendix A.
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from oracle.pyoracle import Reference, ARRAYS  # noqa: E402
from dmlc_core_b200 import datagen  # noqa: E402

GOLD = os.path.join(ROOT, "tests", "golden")

CASES = [
    # name, format, index_bytes, label_column, text
    ("svm_basic", "libsvm", 4, -1, b"1 1:0.5 3:1.25 7:2\n0 2:1e-3 5:-4.5E2\n\n-1 10:3\n"),
    ("svm_crlf", "libsvm", 4, -1, b"1 1:0.5 3:1.25\r\n0 2:1e-3\r\n\r\n-1 10:3\r\n"),
    ("svm_cr_only", "libsvm", 4, -1, b"1 1:0.5\r0 2:2\r"),
    ("svm_qid_binary", "libsvm", 4, -1, b"1 qid:5 1 3\n0 qid:5 2\n1 qid:9 4 6\n"),
    ("svm_qid_values_throws", "libsvm", 4, -1, b"1:0.5 qid:5 1:0.5 3:1.25\n0:1 qid:5 2:1\n"),
    ("svm_weight", "libsvm", 4, -1, b"1:0.25 1:0.5 3:1.25\n0:4 2:1\n"),
    ("svm_weight_qid", "libsvm", 8, -1, b"1:0.25 qid:7 10 20 30\n0:4 qid:7 11\n1:1 qid:8 5 6\n"),
    ("svm_qid_tab", "libsvm", 4, -1, b"1\tqid:3 4\n1:2\tqid:3 4\n1 \t qid:4 5\n1:2  qid:4 5\n"),
    ("svm_chain4", "libsvm", 4, -1, b"1:2:3:4 5:6:7 8\n0 1:2:3:4:5\n"),
    ("svm_spaces_colon", "libsvm", 4, -1, b"1 3 : 0.5 4 :7 5: 8 6:x9\n  0   2:1  \n"),
    ("svm_u32_wrap", "libsvm", 4, -1, b"1 99999999999:1 4294967296:2 4294967295:3\n"),
    ("svm_u64_wrap", "libsvm", 8, -1, b"1 18446744073709551615:1 18446744073709551616:2 99999999999:3\n"),
    ("svm_negative_index_throws", "libsvm", 4, -1, b"1 2:1 -3:0.5\n"),
    ("svm_float_forms", "libsvm", 4, -1,
     b"+1 1:1e5e3 2:1-2 3:.5 4:5. 5:-.5e1 6:1e 7:1e+ 8:- 9:e 10:1e-40 11:1e39 12:4e38\n"
     b"-0 1:0.12345678901234567890123 2:16777217 3:56018346692 4:0.0001e-38 5:1e4294967297\n"),
    ("svm_no_values", "libsvm", 4, -1, b"1 3 5 7\n0 2 4\n1\n0 9\n"),
    ("svm_mixed_values_throws", "libsvm", 4, -1, b"1 3:1 5\n"),
    ("svm_blank_lines_only", "libsvm", 4, -1, b"\n\n  \n\t\n"),
    ("svm_highbytes", "libsvm", 4, -1, b"1 1:2 \xc3\xa9 3:4 \xff5:6\n"),
    ("fm_basic", "libfm", 4, -1, b"1 0:1:0.5 2:3:1.25\n0:2 1:2:3 4:5:6\n-1 7 1:2:1e-2\n"),
    ("fm_novalue", "libfm", 4, -1, b"1 0:1 2:3\n0 7 1:2 9\n"),
    ("fm_mixed_throws", "libfm", 4, -1, b"1 0:1:0.5 2:3\n"),
    ("fm_u64", "libfm", 8, -1, b"1 3:1099511627775:0.5 38:18446744073709551615:2\n"),
    ("fm_chain", "libfm", 4, -1, b"1:2:3:4:5:6:7 8:9:10:11\n0 1:2:3:4\n"),
    ("fm_lone_negative_throws", "libfm", 4, -1, b"1 -5 1:2:3\n"),
    ("csv_basic", "csv", 4, 0, b"1,0.5,2,3\n0,1e-3,-4.5E2,7\n\n-1,3,4,5\n"),
    ("csv_empty_fields", "csv", 4, 0, b"0,,7,\n,,\n1\n,5\n"),
    ("csv_no_label", "csv", 4, -1, b"1,2,3\n4,5,6\n"),
    ("csv_label_last", "csv", 4, 2, b"1,2,3\n4,5,6\n7,8\n"),
    ("csv_crlf", "csv", 4, 1, b"1,2,3\r\n4,5,6\r\n\r\n7,8,9\r\n"),
    ("csv_text_fields", "csv", 4, 0, b"1,abc,2x,nan,inf,0x10,-\n0, 5,\t6,7 \n"),
]

GENERATED = [
    # name, kind, format, index_bytes, label_column, bytes
    ("gen_libsvm", "libsvm", "libsvm", 4, -1, 24 << 10),
    ("gen_csv", "csv", "csv", 4, 0, 24 << 10),
    ("gen_libfm", "libfm", "libfm", 8, -1, 24 << 10),
    ("gen_libsvm_qid", "libsvm_qid", "libsvm", 4, -1, 24 << 10),
    ("gen_criteo", "criteo", "libsvm", 8, -1, 24 << 10),
]

STRTOF = ["1234567901234", "+12345.6789", "-0.00123", "+0123.234e-2", "-234234.123123e20",
          "3.1029831e+38", "000.123e-28", "0", "-0", "-", ".", "e", "1e", "1e+", "1e5e3", "1-2",
          "0.1", "0.3", "1e-40", "1e38", "1e39", "3.5e38", "4e38", "0.0001e-38", "123456789",
          "16777217", "9999999999", "0.12345678901234567890123", "1.00000000000000000000001",
          "1e-8", "1E8", "5.", "-.5", "+.5e1", "340282350000000000000000000000000000000",
          "1e4294967297", "56018346692", "241572670433", "5607233846890", "4194055629409",
          "5968451676", "0." + "0" * 64, "-0." + "0" * 64 + "e3", "0." + "0" * 63 + "1"]
STRTOUINT = ["2147483647", "+12345", "18446744073709551615", "99999999999", "4294967296",
             "-3", "12.5", "", "+-5", "007"]


def main():
    ref = Reference.load()
    if ref is None:
        raise SystemExit("oracle/_ref/libdmlc_ref.so missing: run `make -C oracle ref`")
    os.makedirs(GOLD, exist_ok=True)
    index = []

    def emit(name, fmt, ib, lc, text):
        path = os.path.join(GOLD, name + ".txt")
        with open(path, "wb") as f:
            f.write(text)
        uri = path + ("?format=csv&label_column=%d" % lc if fmt == "csv" else "")
        res = ref.parse_uri(uri, 0, 1, fmt, ib)
        np.savez_compressed(os.path.join(GOLD, name + ".npz"), error=np.int32(res["error"]),
                            **{k: res[k] for k in ARRAYS})
        index.append({"name": name, "format": fmt, "index_bytes": ib, "label_column": lc,
                      "error": int(res["error"]), "rows": int(len(res["label"])),
                      "nnz": int(len(res["index"]))})

    for name, fmt, ib, lc, text in CASES:
        emit(name, fmt, ib, lc, text)
    for name, kind, fmt, ib, lc, nbytes in GENERATED:
        data, _ = datagen.generate(kind, 11, nbytes)
        emit(name, fmt, ib, lc, data.tobytes())
    nums = {"strtof": {t: list(ref.strtof_bits(t)) for t in STRTOF},
            "strtouint32": {t: list(ref.strtouint(t, 4)) for t in STRTOUINT},
            "strtouint64": {t: list(ref.strtouint(t, 8)) for t in STRTOUINT}}
    with open(os.path.join(GOLD, "strtonum.json"), "w") as f:
        json.dump(nums, f, indent=1, sort_keys=True)
    with open(os.path.join(GOLD, "index.json"), "w") as f:
        json.dump(index, f, indent=1)
    for e in index:
        print(e)


if __name__ == "__main__":
    main()
