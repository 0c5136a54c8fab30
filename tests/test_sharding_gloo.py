"""N > 1 host path on CPU: two gloo ranks shard one input by (part_index = rank, num_parts =
world) the way bench.py / a multi-GPU consumer does -- no data-path collective, only the final
bookkeeping is exchanged -- and together must cover the file exactly once."""
import hashlib
import os
import sys

import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _worker(rank, world, port, uri, total, out_dir):
    sys.path.insert(0, ROOT)
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    from dmlc_core_b200 import hostapi
    mine, pieces = hostapi.split_bytes(uri, rank, world, mode=0, chunk_hint=1 << 16, cap=total + 64)
    # each rank owns whole lines only
    assert mine == b"" or mine.endswith(b"\n")
    # bookkeeping a consumer would all-reduce: bytes and rows per shard
    stats = torch.tensor([len(mine), mine.count(b"\n")], dtype=torch.int64)
    gathered = [torch.zeros(2, dtype=torch.int64) for _ in range(world)]
    dist.all_gather(gathered, stats)
    totals = torch.stack(gathered).sum(0)
    with open(os.path.join(out_dir, "part%d.bin" % rank), "wb") as f:
        f.write(mine)
    dist.barrier()
    if rank == 0:
        joined = b"".join(open(os.path.join(out_dir, "part%d.bin" % r), "rb").read() for r in range(world))
        with open(os.path.join(out_dir, "result.txt"), "w") as f:
            f.write("%d %d %s" % (int(totals[0]), int(totals[1]), hashlib.sha1(joined).hexdigest()))
    dist.destroy_process_group()


@pytest.mark.parametrize("world", [2, 3])
def test_two_ranks_cover_the_input_once(tmp_path, world):
    from dmlc_core_b200 import build, datagen
    build.build_cuda()
    build.build_host()
    data, rows = datagen.generate("libsvm_qid", 21, 2 << 20)
    a, b = tmp_path / "a.txt", tmp_path / "b.txt"
    raw = data.tobytes()
    cut = raw.index(b"\n", len(raw) // 2) + 1
    a.write_bytes(raw[:cut])
    b.write_bytes(raw[cut:])
    uri = "%s;%s" % (a, b)
    port = 29500 + (os.getpid() % 2000) + world
    mp.spawn(_worker, args=(world, port, uri, len(raw), str(tmp_path)), nprocs=world, join=True)
    nbytes, nrows, digest = open(tmp_path / "result.txt").read().split()
    assert int(nbytes) == len(raw) and int(nrows) == rows
    assert digest == hashlib.sha1(raw).hexdigest()
