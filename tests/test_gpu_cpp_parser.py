"""GPU suite for the C++ drop-in: dmlc::Parser<IndexType>::Create(uri, part_index, num_parts,
type) -> Next()/Value() over files, against the golden RowBlocks the unmodified reference
produced through the same call, and against the reference itself when oracle/_ref travelled."""
import os
import random

import numpy as np
import pytest

from conftest import GOLDEN, golden_cases, load_golden
from oracle.pyoracle import same_rowblocks

pytestmark = pytest.mark.gpu


@pytest.fixture(scope="module")
def hostapi(cuda_lib):
    from dmlc_core_b200 import build
    build.build_host()
    from dmlc_core_b200 import hostapi as h
    h.load()
    return h


def _trim_hidden_lengths(got, exp):
    """RowBlock hides the true length of weight / qid behind a raw pointer (the reference does
    the same: mixed rows give a shorter array and "no check"). Where the expected
    array is shorter than `size`, only its real entries are comparable through the C++ API."""
    if got.get("error"):
        return got
    rows = len(exp["label"])
    for k in ("weight", "qid"):
        if len(exp[k]) not in (0, rows) and len(got[k]) == rows:
            got[k] = got[k][:len(exp[k])]
    return got


def _uri(path, fmt, lc):
    return path + ("?format=csv&label_column=%d" % lc if fmt == "csv" else "")


@pytest.mark.parametrize("case", golden_cases(), ids=lambda c: c["name"])
def test_create_next_value_matches_golden(hostapi, case):
    _, exp = load_golden(case)
    path = os.path.join(GOLDEN, case["name"] + ".txt")
    uri = _uri(path, case["format"], case["label_column"])
    try:
        p = hostapi.CppParser(uri, 0, 1, case["format"], case["index_bytes"])
        got, _ = p.drain()
        assert p.bytes_read() == os.path.getsize(path)
        p.close()
    except hostapi.DmlcError as e:
        got = {"error": 1, "errmsg": str(e)}
    assert same_rowblocks(_trim_hidden_lengths(got, exp), exp) is None, case["name"]


@pytest.mark.parametrize("kind,fmt,ib,lc", [("libsvm", "libsvm", 4, -1), ("csv", "csv", 4, 0),
                                            ("libfm", "libfm", 8, -1), ("libsvm_qid", "libsvm", 4, -1),
                                            ("criteo", "libsvm", 8, -1)])
def test_sharded_parse_matches_oracle(hostapi, oracle, tmp_path, monkeypatch, kind, fmt, ib, lc):
    """part_index / num_parts sharding (BASELINE configs 4-5): every part equals the oracle on
    that part's bytes; the parts together are the whole file; BytesRead() ends at the part's
    length; a second pass after BeforeFirst() repeats the first; small chunks force many blocks."""
    from dmlc_core_b200 import datagen
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    data, rows = datagen.generate(kind, 8, 5 << 20)
    a, b = tmp_path / "a.txt", tmp_path / "b.txt"
    cut = data.tobytes().index(b"\n", data.size // 3) + 1
    a.write_bytes(data.tobytes()[:cut])
    b.write_bytes(data.tobytes()[cut:])
    uri = _uri(str(a) + ";" + str(b), fmt, lc)
    total_rows = 0
    for nparts in (1, 3):
        labels = []
        for part in range(nparts):
            exp = oracle.parse_files(fmt, [str(a), str(b)], part, nparts, ib, lc)
            p = hostapi.CppParser(uri, part, nparts, fmt, ib)
            got, nblocks = p.drain()
            assert same_rowblocks(got, exp) is None, (kind, nparts, part)
            part_bytes, _, _ = oracle.split_bytes([str(a), str(b)], part, nparts)
            assert p.bytes_read() == len(part_bytes)
            if nparts == 1:
                assert nblocks >= 4
            p.before_first()
            again, _ = p.drain()
            assert same_rowblocks(again, exp) is None
            assert p.bytes_read() == 2 * len(part_bytes)   # not reset, like the reference
            p.close()
            labels.append(got["label"])
        assert sum(len(x) for x in labels) == rows


def test_against_reference_parser_directly(hostapi, reference, tmp_path):
    """Same URI, same (part, nparts) through both Parser::Create implementations."""
    from dmlc_core_b200 import datagen
    for kind, fmt, ib, lc in (("libsvm", "libsvm", 4, -1), ("csv", "csv", 4, 0), ("libfm", "libfm", 4, -1)):
        data, _ = datagen.generate(kind, 12, 2 << 20)
        f = tmp_path / (kind + ".txt")
        f.write_bytes(data.tobytes())
        uri = _uri(str(f), fmt, lc)
        for part, nparts in ((0, 1), (1, 2), (3, 4)):
            ref = reference.parse_uri(uri, part, nparts, fmt, ib)
            p = hostapi.CppParser(uri, part, nparts, fmt, ib)
            got, _ = p.drain()
            p.close()
            assert same_rowblocks(got, ref) is None, (kind, part, nparts)


def test_row_block_iter_numcol(hostapi, oracle, tmp_path):
    from dmlc_core_b200 import datagen
    data, rows = datagen.generate("libsvm", 4, 1 << 20)
    f = tmp_path / "it.svm"
    f.write_bytes(data.tobytes())
    exp = oracle.parse_buffer("libsvm", data.tobytes(), 4)
    got, ncol, _ = hostapi.row_iter(str(f), 0, 1, "libsvm", 4)
    assert same_rowblocks(got, exp) is None
    assert ncol == int(exp["index"].max()) + 1


def test_long_line_and_error_propagation(hostapi, oracle, tmp_path, monkeypatch):
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    # one line of ~3 MiB (longer than the 1 MiB staging buffer), then normal lines
    line = b"1 " + b" ".join(b"%d:0.5" % i for i in range(1, 330000)) + b"\n"
    text = line + b"0 1:2 3:4\n" * 1000
    f = tmp_path / "long.svm"
    f.write_bytes(text)
    exp = oracle.parse_buffer("libsvm", text, 4)
    p = hostapi.CppParser(str(f), 0, 1, "libsvm", 4)
    got, _ = p.drain()
    p.close()
    assert same_rowblocks(got, exp) is None
    bad = tmp_path / "bad.svm"
    bad.write_bytes(b"1 1:1\n0 -2:1\n")
    p = hostapi.CppParser(str(bad), 0, 1, "libsvm", 4)
    with pytest.raises(hostapi.DmlcError, match="sign == true"):
        p.drain()
    p.close()


def test_rewind_after_a_parse_error_starts_from_the_top(hostapi, tmp_path, monkeypatch):
    """Two chunks that each hold a negative index are in flight when Next() throws for the first;
    BeforeFirst() must drop both, so that the next pass begins with the file's first block again
    and not with what was queued behind the failure."""
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    monkeypatch.setenv("DMLC_B200_DEPTH", "3")
    good = b"".join(b"%d 1:%d 2:0.5\n" % (i % 2, i) for i in range(120000))     # ~1.7 MiB
    bad = b"0 -2:1\n"
    f = tmp_path / "twice_bad.svm"
    f.write_bytes(good[:1 << 20].rsplit(b"\n", 1)[0] + b"\n" + bad + good + bad + good)
    fresh = hostapi.CppParser(str(f), 0, 1, "libsvm", 4)
    first = fresh.next()
    fresh.close()
    p = hostapi.CppParser(str(f), 0, 1, "libsvm", 4)
    for _ in range(3):       # each pass: the first block, then the failure
        b = p.next()
        assert b is not None and np.array_equal(b["index"], first["index"])
        assert np.array_equal(b["value"].view(np.uint32), first["value"].view(np.uint32))
        with pytest.raises(hostapi.DmlcError, match="sign == true"):
            while p.next() is not None:
                pass
        p.before_first()
    p.close()


def test_true_dropin_build_against_reference_headers(hostapi):
    """oracle/_ref/libdmlc_dropin.so = the reference's headers, InputSplit, filesystem and
    registry with src/data.cc replaced by this package's parser.
    Built only where /root/reference exists; it travels to the GPU box like libdmlc_ref.so."""
    from conftest import ROOT
    path = os.path.join(ROOT, "oracle", "_ref", "libdmlc_dropin.so")
    if not os.path.exists(path):
        pytest.skip("drop-in build not present")
    lib = hostapi.load(path)
    for case in golden_cases():
        _, exp = load_golden(case)
        f = os.path.join(GOLDEN, case["name"] + ".txt")
        try:
            p = hostapi.CppParser(_uri(f, case["format"], case["label_column"]), 0, 1,
                                  case["format"], case["index_bytes"], lib=lib)
            got, _ = p.drain()
            assert p.bytes_read() == os.path.getsize(f)
            p.close()
        except hostapi.DmlcError as e:
            got = {"error": 1, "errmsg": str(e)}
        assert same_rowblocks(_trim_hidden_lengths(got, exp), exp) is None, case["name"]


def test_disk_cache_round_trip_with_the_reference(hostapi, reference, oracle, tmp_path, monkeypatch):
    """"uri#cachefile" (SURVEY 8f-3): pages written here in the reference's RowBlockContainer::Save
    layout are read back by the reference's own DiskRowIter, and a second Create here replays the
    cache without parsing."""
    from dmlc_core_b200 import datagen
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    # csv: label_column=0 in the URI does not reach the parser through RowBlockIter::Create
    # (reference src/data.cc:95-97), so the expectation is "no label column"
    for kind, fmt, ib, lc in (("libsvm", "libsvm", 4, -1), ("libfm", "libfm", 8, -1), ("csv", "csv", 4, -1)):
        data, rows = datagen.generate(kind, 15, 3 << 20)
        f = tmp_path / (kind + ".txt")
        f.write_bytes(data.tobytes())
        cache = tmp_path / (kind + ".cache")
        args = "?format=csv&label_column=0" if fmt == "csv" else ""
        uri = "%s%s#%s" % (f, args, cache)
        exp = oracle.parse_buffer(fmt, data.tobytes(), ib, lc)
        got, ncol, nblocks = hostapi.row_iter(uri, 0, 1, fmt, ib)          # parses on the GPU, writes pages
        assert same_rowblocks(got, exp) is None
        assert ncol == int(exp["index"].max()) + 1
        size = os.path.getsize(cache)
        theirs, _ = reference.iter_uri(uri, 0, 1, fmt, ib)                  # the reference replays OUR cache
        assert same_rowblocks(theirs, exp) is None
        f.write_bytes(b"")                                                   # text gone: must come from the cache
        again, ncol2, _ = hostapi.row_iter(uri, 0, 1, fmt, ib, passes=2)
        assert same_rowblocks(again, exp) is None and ncol2 == ncol
        assert os.path.getsize(cache) == size


def test_two_parsers_interleaved_in_one_process(hostapi, oracle, tmp_path, monkeypatch):
    """One process, one parser per part (DMLC_B200_DEVICE=auto spreads them over the box's GPUs;
    with a single GPU both land on device 0): handles are independent, calls may interleave."""
    from dmlc_core_b200 import datagen
    monkeypatch.setenv("DMLC_B200_DEVICE", "auto")
    monkeypatch.setenv("DMLC_B200_NUM_DEVICES", "1")
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    data, rows = datagen.generate("libfm", 23, 4 << 20)
    f = tmp_path / "fm.txt"
    f.write_bytes(data.tobytes())
    parsers = [hostapi.CppParser(str(f), k, 2, "libfm", 8) for k in range(2)]
    blocks = [[], []]
    live = [True, True]
    while any(live):
        for k in range(2):
            if live[k]:
                b = parsers[k].next()
                if b is None:
                    live[k] = False
                else:
                    blocks[k].append(b)
    for k in range(2):
        exp = oracle.parse_files("libfm", [str(f)], k, 2, 8)
        got = hostapi.concat_blocks(blocks[k], 8)
        assert same_rowblocks(got, exp) is None, k
        parsers[k].close()


def test_rewind_and_destroy_while_the_feeder_is_running(hostapi, oracle, tmp_path, monkeypatch):
    """BeforeFirst() and delete in the middle of a pass: the feeder thread is stopped with chunks
    in flight and a read under way, what was in flight is dropped, and the next pass starts from
    the first row again (reference text_parser.h:40-42 / threadediter.h BeforeFirst semantics)."""
    from dmlc_core_b200 import datagen
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    data, rows = datagen.generate("libsvm", 31, 12 << 20)
    f = tmp_path / "big.svm"
    f.write_bytes(data.tobytes())
    exp = oracle.parse_files("libsvm", [str(f)], 0, 1, 4)
    p = hostapi.CppParser(str(f), 0, 1, "libsvm", 4)
    for taken in (0, 1, 3, 5):
        for _ in range(taken):
            assert p.next() is not None
        p.before_first()
    got, nblocks = p.drain()
    assert nblocks >= 8 and same_rowblocks(got, exp) is None
    assert p.next() is None and p.next() is None      # stays at the end
    p.before_first()
    assert p.next() is not None
    p.close()                                         # destroyed mid-pass
    for _ in range(3):                                # created and dropped without a single Next()
        hostapi.CppParser(str(f), 0, 1, "libsvm", 4).close()


def test_row_block_iter_drops_uri_arguments_like_the_reference(hostapi, oracle, tmp_path):
    """Reference quirk kept on purpose: CreateIter_ hands CreateParser_ the URI *without* its
    "?key=value" part (src/data.cc:95-97), so through RowBlockIter::Create a CSV file never has a
    label column and type "auto" never sees format=.  Parser::Create keeps the arguments."""
    f = tmp_path / "t.csv"
    rnd = random.Random(5)
    f.write_text("".join(",".join("%.3f" % rnd.uniform(-5, 5) for _ in range(20)) + "\n" for _ in range(300)))
    uri = str(f) + "?format=csv&label_column=0"
    got, ncol, _ = hostapi.row_iter(uri, 0, 1, "csv", 4)
    exp = oracle.parse_buffer("csv", f.read_bytes(), 4, -1)          # no label column
    assert same_rowblocks(got, exp) is None and ncol == 20
    assert not got["label"].any()
    with_args, _ = hostapi.CppParser(uri, 0, 1, "csv", 4).drain()    # Parser::Create honours them
    assert same_rowblocks(with_args, oracle.parse_buffer("csv", f.read_bytes(), 4, 0)) is None
    g = tmp_path / "positive.csv"     # read as LibSVM a '-' would be a negative index: keep it parseable
    g.write_text("".join(",".join("%.3f" % rnd.uniform(0.5, 5) for _ in range(20)) + "\n" for _ in range(300)))
    uri_g = str(g) + "?format=csv&label_column=0"
    as_svm, _, _ = hostapi.row_iter(uri_g, 0, 1, "auto", 4)          # "auto" -> libsvm, format= lost
    assert same_rowblocks(as_svm, oracle.parse_buffer("libsvm", g.read_bytes(), 4)) is None
    from oracle.pyoracle import Reference
    reference = Reference.load()       # present on the GPU box when oracle/_ref travelled along
    if reference is not None:
        theirs, their_ncol = reference.iter_uri(uri, 0, 1, "csv", 4)
        assert same_rowblocks(got, theirs) is None and their_ncol == ncol
        theirs, _ = reference.iter_uri(uri_g, 0, 1, "auto", 4)
        assert same_rowblocks(as_svm, theirs) is None


def test_parser_on_stdin(hostapi, oracle, tmp_path):
    """Parser::Create("stdin", 0, 1, type): the `cat file | program stdin` use of the reference
    (SingleFileSplit behind InputSplit::Create, src/io.cc:96-98).  The feeder thread reads the
    pipe straight into the pinned staging buffers; result equals the oracle on the same bytes."""
    import subprocess
    import sys
    from dmlc_core_b200 import datagen
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    data, rows = datagen.generate("libsvm", 41, 9 << 20)
    exp = oracle.parse_buffer("libsvm", data.tobytes(), 4)
    code = ("import sys, numpy as np; sys.path.insert(0, %r); from dmlc_core_b200 import hostapi; "
            "got, n = hostapi.CppParser('stdin', 0, 1, 'libsvm', 4).drain(); "
            "np.savez(sys.argv[1], nblocks=n, **{k: v for k, v in got.items() if hasattr(v, 'dtype')})" % root)
    out = tmp_path / "got.npz"
    env = dict(os.environ, DMLC_B200_CHUNK_MB="1")
    subprocess.run([sys.executable, "-c", code, str(out)], input=data.tobytes(), check=True, env=env)
    z = np.load(out)
    got = {k: z[k] for k in ("offset", "label", "weight", "qid", "field", "index", "value")}
    got["error"] = 0
    assert same_rowblocks(got, exp) is None and int(z["nblocks"]) >= 6


@pytest.mark.parametrize("mode", ["in place", "in place, not pinned", "copying"])
def test_row_block_iter_modes_agree(hostapi, oracle, tmp_path, monkeypatch, mode):
    """RowBlockIter::Create lets blocks land in its own (pinned) arrays; on a host that refuses to
    pin, and with the copying path forced, the result is the same."""
    from dmlc_core_b200 import datagen
    monkeypatch.setenv("DMLC_B200_CHUNK_MB", "1")
    if mode == "in place, not pinned":
        monkeypatch.setenv("DMLC_B200_ITER_NO_PIN", "1")
    if mode == "copying":
        monkeypatch.setenv("DMLC_B200_ITER_COPY", "1")
    for kind, fmt, ib in (("libsvm_qid", "libsvm", 4), ("libfm", "libfm", 8), ("csv", "csv", 4)):
        data, rows = datagen.generate(kind, 19, 7 << 20)
        f = tmp_path / (kind + ".txt")
        f.write_bytes(data.tobytes())
        exp = oracle.parse_buffer(fmt, data.tobytes(), ib, -1)
        got, ncol, _ = hostapi.row_iter(str(f), 0, 1, fmt, ib)
        assert same_rowblocks(_trim_hidden_lengths(got, exp), exp) is None, (mode, kind)
        assert ncol == int(exp["index"].max()) + 1


@pytest.mark.parametrize("kind,fmt,ib", [("libfm", "libfm", 8), ("csv", "csv", 4)])
def test_a_gigabyte_through_parser_create_matches_the_reference(hostapi, reference, tmp_path, kind, fmt, ib):
    """Full-size check of the drop-in API: 1 GB of text in one file through Parser<I>::Create here
    and through the reference's own Parser<I>::Create, compared by a streamed FNV-1a of each of the
    seven arrays (row lengths instead of offsets, so block granularity does not matter)."""
    from dmlc_core_b200 import datagen
    shm = "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else str(tmp_path)
    data, rows = datagen.generate(kind, 77, 250 * 1000 * 1000)
    path = os.path.join(shm, "b200_test_gigabyte_%s.txt" % kind)
    try:
        with open(path, "wb") as f:
            for _ in range(4):
                data.tofile(f)
        uri = path + ("?format=csv&label_column=0" if fmt == "csv" else "")
        for part, nparts in ((0, 1), (1, 3)):
            mine = hostapi.time_parser(uri, part, nparts, fmt, ib, passes=1, want_hash=True)
            theirs, ref_rows, ref_nnz = reference.hash_uri(uri, part, nparts, fmt, ib)
            assert mine["rows"] == ref_rows and mine["nnz"] == ref_nnz, (part, nparts)
            assert mine["hash"] == theirs, (part, nparts)
            if nparts == 1:
                # BytesRead() keeps counting across BeforeFirst (text_parser.h:101, never reset): hash pass + timed pass
                assert ref_rows == 4 * rows and mine["bytes_read"] == 2 * 4 * data.size
    finally:
        if os.path.exists(path):
            os.remove(path)
