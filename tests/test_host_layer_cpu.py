"""CPU suite for the C++ host layer (dmlc_core_b200/src): InputSplit partition arithmetic against
the compiled reference, URI handling and error behaviour.  No GPU work happens here."""
import os
import random

import pytest


@pytest.fixture(scope="module")
def hostapi():
    from dmlc_core_b200 import build
    build.build_cuda()
    build.build_host()
    from dmlc_core_b200 import hostapi as h
    h.load()
    return h


def _files(tmp_path, rnd, nfiles, eol_choices):
    paths = []
    for i in range(nfiles):
        lines = []
        for _ in range(rnd.randint(3, 80)):
            lines.append(" ".join("%d:%d" % (rnd.randint(1, 999), rnd.randint(0, 9))
                                  for _ in range(rnd.randint(0, 12))) + rnd.choice(eol_choices))
        p = tmp_path / ("f%02d.txt" % i)
        p.write_bytes(("1 " + "0 ".join(lines)).encode())
        paths.append(str(p))
    return paths


def test_partitions_match_reference(hostapi, reference, tmp_path):
    """InputSplit::Create(uri, k, n, "text") yields exactly the bytes the reference yields
    (ResetPartition / SeekRecordBegin semantics), for single files, ';' lists, CRLF, blank lines,
    tiny chunk hints, more parts than lines."""
    rnd = random.Random(17)
    for trial, (nfiles, eols) in enumerate([(1, ["\n"]), (3, ["\n", "\r\n", "\n\n"]), (4, ["\r"]),
                                            (2, ["\n", "\n\n\n"])]):
        d = tmp_path / ("t%d" % trial)
        d.mkdir()
        paths = _files(d, rnd, nfiles, eols)
        uri = ";".join(paths)
        total = sum(os.path.getsize(p) for p in paths)
        for nparts in (1, 2, 3, 7, 16, 200):
            joined = b""
            for part in range(nparts):
                want, n, _ = reference.split_bytes(uri, part, nparts, total + 64)
                got, _ = hostapi.split_bytes(uri, part, nparts, mode=0, cap=total + 64)
                assert got == want, (trial, nparts, part)
                joined += got
            assert len(joined) == total


def test_chunks_are_line_aligned_and_records_complete(hostapi, tmp_path):
    from dmlc_core_b200 import datagen
    data, rows = datagen.generate("libsvm", 9, 9 << 20)
    p = tmp_path / "big.svm"
    p.write_bytes(data.tobytes())
    got, pieces = hostapi.split_bytes(str(p), 0, 1, mode=0, cap=data.size + 64)
    assert got == data.tobytes() and pieces >= 2       # default 8 MiB chunks
    recs, nrec = hostapi.split_bytes(str(p), 1, 3, mode=1, cap=data.size + 64)
    part, _ = hostapi.split_bytes(str(p), 1, 3, mode=0, cap=data.size + 64)
    # (the last record of a chunk keeps its EOL byte, as in the reference's ExtractNextRecord)
    assert [r for r in recs.split(b"\n") if r] == [r for r in part.split(b"\n") if r]
    assert nrec == part.count(b"\n")


def test_directory_and_regex_uris(hostapi, tmp_path):
    d = tmp_path / "dir"
    d.mkdir()
    for name, body in (("a.svm", b"1 1:1\n"), ("b.svm", b"0 2:2\n"), ("c.txt", b"1 3:3\n"), ("empty.svm", b"")):
        (d / name).write_bytes(body)
    whole, _ = hostapi.split_bytes(str(d), 0, 1)
    assert sorted(whole.split(b"\n")[:-1]) == [b"0 2:2", b"1 1:1", b"1 3:3"]
    only, _ = hostapi.split_bytes(str(d) + "/.*\\.svm", 0, 1)
    assert sorted(only.split(b"\n")[:-1]) == [b"0 2:2", b"1 1:1"]
    with pytest.raises(hostapi.DmlcError):
        hostapi.split_bytes(str(d) + "/nothing_here", 0, 1)
    with pytest.raises(hostapi.DmlcError):
        hostapi.split_bytes(str(d / "a.svm"), 3, 3)        # part_index must be < num_parts


def test_create_errors_match_reference_behaviour(hostapi, tmp_path):
    """Unknown types, csv with a 64-bit index, bad csv arguments: all raise (reference
    src/data.cc:80-82,158; csv_parser.h:50-51). On a box without a GPU a valid Create raises too
    -- the product has no CPU path."""
    import torch
    f = tmp_path / "x.csv"
    f.write_bytes(b"1,2,3\n")
    with pytest.raises(hostapi.DmlcError, match="Unknown data type"):
        hostapi.CppParser(str(f), 0, 1, "parquet")
    with pytest.raises(hostapi.DmlcError, match="Unknown data type csv"):
        hostapi.CppParser(str(f) + "?format=csv", 0, 1, "auto", index_bytes=8)
    with pytest.raises(hostapi.DmlcError, match="Cannot find argument"):
        hostapi.CppParser(str(f) + "?format=csv&delimiter=;", 0, 1, "csv")
    with pytest.raises(hostapi.DmlcError):
        hostapi.CppParser(str(f) + "?format=csv&label_column=x", 0, 1, "csv")
    with pytest.raises(hostapi.DmlcError):
        hostapi.CppParser(str(f) + "#a#b", 0, 1, "csv")
    if not torch.cuda.is_available():
        with pytest.raises(hostapi.DmlcError, match="no CPU path|cannot create"):
            hostapi.CppParser(str(f) + "?format=csv&label_column=0", 0, 1, "auto")


def test_cache_written_by_the_reference_loads_here(hostapi, reference, oracle, tmp_path):
    """"uri#cachefile": a page cache built by the reference's DiskRowIter
    (src/data/disk_row_iter.h:111-141, RowBlockContainer::Save row_block.h:189-201) is read back by
    this package's RowBlockIter without parsing anything (no GPU involved)."""
    from dmlc_core_b200 import datagen
    from oracle.pyoracle import same_rowblocks
    for kind, fmt, ib in (("libsvm", "libsvm", 4), ("libfm", "libfm", 8), ("libsvm_qid", "libsvm", 4)):
        data, rows = datagen.generate(kind, 14, 600 << 10)
        f = tmp_path / (kind + ".txt")
        f.write_bytes(data.tobytes())
        cache = tmp_path / (kind + ".cache")
        uri = "%s#%s" % (f, cache)
        ref_blocks, ref_ncol = reference.iter_uri(uri, 0, 1, fmt, ib)     # builds the cache
        assert os.path.getsize(cache) > 0 and ref_blocks["error"] == 0
        got, ncol, _ = hostapi.row_iter(uri, 0, 1, fmt, ib, passes=2)
        exp = oracle.parse_buffer(fmt, data.tobytes(), ib)
        assert same_rowblocks(got, exp) is None
        assert ncol == int(exp["index"].max()) + 1


def test_stdin_source(hostapi, tmp_path):
    """InputSplit::Create("stdin", ...) (reference src/io.cc:96-98 -> SingleFileSplit): the lines
    arriving on descriptor 0, through a pipe and from a redirected file, as chunks of whole lines
    and as records; a line longer than the chunk hint grows the chunk."""
    import subprocess
    import sys
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    rnd = random.Random(9)
    lines = [" ".join("%d:%d" % (rnd.randint(1, 99999), rnd.randint(0, 9)) for _ in range(rnd.randint(0, 40)))
             for _ in range(20000)]
    lines[700] = "7 " + "1:1 " * 200000            # ~800 KB line, longer than the 256 KiB minimum chunk
    text = ("\n".join(lines) + "\n").encode()
    code = ("import sys; sys.path.insert(0, %r); from dmlc_core_b200 import hostapi; "
            "b, n = hostapi.split_bytes('stdin', 0, 1, mode=int(sys.argv[1]), chunk_hint=1 << 18); "
            "sys.stdout.buffer.write(b); sys.stderr.write(str(n))" % root)
    for mode in (0, 1):
        piped = subprocess.run([sys.executable, "-c", code, str(mode)], input=text, capture_output=True, check=True)
        pieces = int(piped.stderr.decode().strip().split()[-1])
        if mode == 0:
            assert piped.stdout == text and pieces > 3
        else:
            # a record keeps the EOL bytes the reference's ExtractNextRecord leaves in it (all but
            # the last of its run, or all of them at the end of a chunk): compare line by line
            kept = [x.encode() for x in lines if x]
            assert [x for x in piped.stdout.split(b"\n") if x] == kept and pieces == len(kept)
    f = tmp_path / "in.txt"
    f.write_bytes(text[:-1])                         # last line without an EOL
    with open(f, "rb") as fh:
        redirected = subprocess.run([sys.executable, "-c", code, "0"], stdin=fh, capture_output=True, check=True)
    assert redirected.stdout == text[:-1]


def test_large_chunks_go_through_the_reader_pool(hostapi, tmp_path, monkeypatch):
    """Chunks of several MiB are read by the persistent reader pool (TextSplit::ReadRange fans a
    chunk out in >= 2 MiB pieces): the bytes must come back complete and in order, for every
    partition, also when the pool is rebuilt for another thread count."""
    rnd = random.Random(77)
    line = lambda: " ".join("%d:%.4f" % (rnd.randint(1, 10 ** 6), rnd.random()) for _ in range(rnd.randint(1, 30)))
    text = ("\n".join(line() for _ in range(90000)) + "\n").encode()
    assert len(text) > 18 << 20
    f = tmp_path / "big.txt"
    f.write_bytes(text)
    for threads, use_map in (("3", "1"), ("1", "1"), ("7", "0"), ("2", "0")):   # mmap + streaming stores / pread
        monkeypatch.setenv("DMLC_B200_READ_THREADS", threads)
        monkeypatch.setenv("DMLC_B200_READ_MMAP", use_map)
        whole, pieces = hostapi.split_bytes(str(f), 0, 1, mode=0, chunk_hint=9 << 20, cap=len(text) + 64)
        assert whole == text and 2 <= pieces <= 4
    parts = [hostapi.split_bytes(str(f), k, 3, mode=0, chunk_hint=5 << 20, cap=len(text) + 64)[0] for k in range(3)]
    assert b"".join(parts) == text and all(parts)


def test_input_split_cache_file_is_interchangeable_with_the_reference(hostapi, reference, tmp_path):
    """InputSplit::Create("uri#cachefile", ...) (reference CachedInputSplit,
    src/io/cached_input_split.h): the first pass records the chunks, later passes and later
    processes replay them without the original files; caches written by either implementation
    are read by the other."""
    rnd = random.Random(21)
    paths = _files(tmp_path, rnd, 3, ["\n", "\r\n", "\n"])
    uri = ";".join(paths)
    total = sum(os.path.getsize(p) for p in paths)
    for nparts in (1, 2):
        for part in range(nparts):
            want, _, _ = reference.split_bytes(uri, part, nparts, total + 64)
            mine = tmp_path / ("mine%d" % nparts)
            theirs = tmp_path / ("theirs%d" % nparts)
            got, _ = hostapi.split_bytes("%s#%s" % (uri, mine), part, nparts, cap=total + 64)     # records
            assert got == want
            again, _ = hostapi.split_bytes("%s#%s" % (uri, mine), part, nparts, cap=total + 64)   # replays
            assert again == want
            ref_reads_mine, _, _ = reference.split_bytes("%s#%s" % (uri, mine), part, nparts, total + 64)
            assert ref_reads_mine == want
            ref_first, _, _ = reference.split_bytes("%s#%s" % (uri, theirs), part, nparts, total + 64)  # records
            assert ref_first == want
            i_read_theirs, _ = hostapi.split_bytes("%s#%s" % (uri, theirs), part, nparts, mode=1, cap=total + 64)
            lines = lambda b: [x for x in b.replace(b"\r", b"\n").split(b"\n") if x]
            assert lines(i_read_theirs) == lines(want)
    decorated = sorted(p.name for p in tmp_path.iterdir() if p.name.startswith("mine2"))
    assert decorated == ["mine2.split2.part0", "mine2.split2.part1"]      # URISpec's suffix, uri_spec.h:59-66


def test_input_split_shuffle_matches_reference(reference, tmp_path):
    """InputSplitShuffle of the stand-alone layer: for a fixed seed the sub-partitions come out in
    the reference's order, on the first pass and after every BeforeFirst (which reshuffles)."""
    import ctypes as C
    from dmlc_core_b200 import build, hostapi
    build.build_host()
    lib = hostapi.load()
    lib.dmlcb_shuffle_bytes.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                        C.c_char_p, C.c_uint64, C.c_char_p, C.c_uint64]
    lib.dmlcb_shuffle_bytes.restype = C.c_uint64
    ref = reference.lib
    ref.refshim_shuffle_bytes.argtypes = [C.c_char_p, C.c_uint, C.c_uint, C.c_uint, C.c_int, C.c_int,
                                          C.c_char_p, C.c_uint64]
    ref.refshim_shuffle_bytes.restype = C.c_uint64
    rnd = random.Random(4)
    files = []
    for k in range(3):
        f = tmp_path / ("part%d.txt" % k)
        f.write_bytes(b"".join(b"%d %d:%d\n" % (k, i, rnd.randint(0, 10 ** rnd.randint(1, 9)))
                               for i in range(rnd.randint(200, 900))))
        files.append(str(f))
    uri = ";".join(files).encode()
    cap = 1 << 20
    for nparts, shuffle_parts, seed in ((1, 1, 0), (1, 4, 0), (2, 3, 7), (3, 5, 123), (2, 8, 1)):
        for part in range(nparts):
            mine, theirs = C.create_string_buffer(cap), C.create_string_buffer(cap)
            err = C.create_string_buffer(512)
            n1 = lib.dmlcb_shuffle_bytes(uri, part, nparts, shuffle_parts, seed, 3, mine, cap, err, 512)
            n2 = ref.refshim_shuffle_bytes(uri, part, nparts, shuffle_parts, seed, 3, theirs, cap)
            assert n1 == n2 and n1 < cap, (err.value, n1, n2)
            assert mine.raw[:n1] == theirs.raw[:n2], (nparts, shuffle_parts, seed, part)
