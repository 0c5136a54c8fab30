#!/usr/bin/env python
"""bench.py -- parsed input GB/s of the B200 text -> CSR path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Headline workload (N=1): BASELINE.json configs[2], the largest single-GPU configuration -- LibFM
text with field:index:value triples, uint64 indices, 20 GB synthetic, Parser<uint64_t>, one B200.
One step = one pass of the hot path over the whole 20 GB (one seeded 1 000 MB block tiled 20x in
HBM).  With N > 1 (torchrun, one rank per GPU) every rank parses its own 20 GB with no collective
on the data path ("weak" scaling); in addition ONE shared file is parsed by all ranks together,
rank g taking Parser::Create(uri, g, N, type), the InputSplit part_index sharding of the reference.

Printed JSON (rank 0):
  value     whole-job GB/s of input text, text already resident in HBM (CUDA events)
  e2e       same metric through the C ABI (b200p_submit / b200p_next) with the text in pinned HOST
            memory: H2D of every chunk and D2H of every CSR array inside the timed region
  e2e_device_out     the same with out_residency = device (CSR stays in HBM for a GPU consumer)
  e2e_parser_create  dmlc::Parser<I>::Create(uri, part, nparts, type) of libdmlc_b200_host.so on a
            file in /dev/shm: the same API on the same kind of source as the reference arm
  parity_in_bench    FNV-1a of the seven arrays of a sample of the timed text, ours == reference's
  roofline  the dominant kernel (ParseTiles) against the measured HBM copy peak
  shapes    all five BASELINE configs (N=1): per-kernel ms, value, e2e, roofline, reference CPU rate
  cpu_baseline  the unmodified reference parser (oracle/_ref) on this box's host cores
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

import numpy as np  # noqa: E402

GB = 1e9

# BASELINE.json configs, in order.  gb = the size the config names (per GPU for the sharded ones).
SHAPES = [
    {"key": "configs[0]", "what": "LibSVM, ~30 nnz/row, Parser<uint32_t> (100 MB on CPU in libsvm_parser_test)",
     "kind": "libsvm", "fmt": "libsvm", "ib": 4, "lc": -1, "gb": 0.1},
    {"key": "configs[1]", "what": "dense CSV, 128 float columns + label_column=0, Parser<uint32_t>, 10 GB",
     "kind": "csv", "fmt": "csv", "ib": 4, "lc": 0, "gb": 10.0},
    {"key": "configs[2]", "what": "LibFM field:index:value triples, uint64 indices, Parser<uint64_t>, 20 GB",
     "kind": "libfm", "fmt": "libfm", "ib": 8, "lc": -1, "gb": 20.0},
    {"key": "configs[3]", "what": "LibSVM with qid and weights (value-less features), Parser<uint32_t>, 50 GB sharded by part_index",
     "kind": "libsvm_qid", "fmt": "libsvm", "ib": 4, "lc": -1, "gb": 50.0},
    {"key": "configs[4]", "what": "Criteo-shaped LibSVM (13 dense + 26 hashed), uint64 indices, 200 GB across 8 GPUs",
     "kind": "criteo", "fmt": "libsvm", "ib": 8, "lc": -1, "gb": 25.0},
]
HEADLINE = 2


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--shape", type=int, default=HEADLINE, help="index into BASELINE.json configs (headline)")
    ap.add_argument("--gb", type=float, default=0.0, help="text bytes per GPU per step (GB); 0 = what the config names, at most 20")
    ap.add_argument("--block-mb", type=int, default=1000, help="unique generated block (MB)")
    ap.add_argument("--chunk-mb", type=int, default=64, help="e2e chunk size (MiB)")
    ap.add_argument("--range-bytes", type=int, default=0)
    ap.add_argument("--depth", type=int, default=3, help="e2e chunks in flight")
    ap.add_argument("--streams", type=int, default=2,
                    help="device-resident measurement: handles/streams the windows of a step alternate between")
    ap.add_argument("--host-threads", type=int, default=-1,
                    help="e2e: b200p_set_host_threads (-1 = library default, 0 = copy every array back)")
    ap.add_argument("--cpu-seconds", type=float, default=10.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    ap.add_argument("--no-shapes", action="store_true")
    ap.add_argument("--shape-mb", type=int, default=1000, help="block size of the per-shape measurements (MB)")
    ap.add_argument("--file-gb", type=float, default=4.0, help="size of the /dev/shm file behind e2e_parser_create (GB, per rank)")
    return ap.parse_args()


def ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the
    committed `ncu --set full` capture of this same workload (profiles/current_ncu.metrics.json,
    written by profiles/summarize_ncu.py). bench.py itself never runs under a profiler."""
    try:
        with open(os.path.join(ROOT, "profiles", "current_ncu.metrics.json")) as f:
            for k in json.load(f):
                if kernel_substr in k.get("kernel", ""):
                    return {"bytes_per_launch": float(k["dram_bytes_per_launch"]),
                            "captured_on_bytes": k.get("text_bytes"), "capture": k.get("capture")}
    except Exception:
        pass
    return None


def measured_peak():
    try:
        with open(os.path.join(ROOT, "MEASURED_PEAKS.json")) as f:
            return float(json.load(f)["hbm_gbs"]), "measured (MEASURED_PEAKS.json)"
    except Exception:
        return 6650.0, "fallback (B200_PROFILING.md)"


class ClockSampler(object):
    """nvidia-smi clocks / throttle reasons sampled during the timed region."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, device):
        self.device = device
        self.proc = None
        self.lines = []

    def start(self):
        try:
            self.proc = subprocess.Popen(
                ["nvidia-smi", "-i", str(self.device), "--query-gpu=" + self.Q,
                 "--format=csv,noheader,nounits", "-lms", "20"],
                stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
            self.thread = threading.Thread(target=self._pump, daemon=True)
            self.thread.start()
        except Exception:
            self.proc = None

    def _pump(self):
        for line in self.proc.stdout:
            self.lines.append(line.strip())

    def stop(self):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["nvidia-smi unavailable"]}
        time.sleep(0.15)
        self.proc.terminate()
        try:
            self.proc.wait(timeout=2)
        except Exception:
            self.proc.kill()
        sm, mx, reasons, pw = [], [], set(), []
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        for ln in self.lines:
            f = [x.strip() for x in ln.split(",")]
            if len(f) < 9:
                continue
            try:
                sm.append(float(f[1]))
                mx.append(float(f[2]))
                pw.append(float(f[3]))
            except ValueError:
                continue
            for name, val in zip(names, f[5:9]):
                if val.lower().startswith("active"):
                    reasons.add(name)
        if not sm:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["no samples"]}
        return {"sm_mhz": float(np.median(sm)), "sm_max_mhz": float(max(mx)),
                "power_w_max": float(max(pw)), "samples": len(sm), "reasons": sorted(reasons)}


def make_block(shape, block_mb, pinned, seed=1):
    """One unique block of the shape's text (whole lines), padded with '\\n' to a multiple of 16
    bytes (blank lines are skipped by the reference parsers) so tiled copies stay aligned."""
    import torch
    from dmlc_core_b200 import datagen
    target = block_mb * 1000 * 1000
    cap = target + (8 << 20)
    host = torch.empty(cap, dtype=torch.uint8, pin_memory=pinned)
    arr = host.numpy()
    kw = {"ncols": 128} if shape["kind"] == "csv" else {}
    data, rows = datagen.generate(shape["kind"], seed, target, out=arr, **kw)
    n = data.size
    pad = (-n) % 16
    arr[n:n + pad] = 10
    return host, n + pad, rows


def cut_chunks(arr, n, chunk_bytes):
    """[(begin, end)] chunks of <= chunk_bytes cut after a newline (what InputSplit hands out,
    reference src/io/input_split_base.cc:211-239)."""
    cuts, pos = [], 0
    while pos < n:
        end = min(n, pos + chunk_bytes)
        if end < n:
            back = arr[pos:end][::-1]
            k = int(np.argmax(back == 10))
            end -= k
        cuts.append((pos, end))
        pos = end
    return cuts


def chunks_end_at(arr, limit):
    """largest prefix <= limit that ends right after a newline"""
    back = arr[:limit][::-1]
    return limit - int(np.argmax(back == 10))


def shm_dir():
    return "/dev/shm" if os.path.isdir("/dev/shm") and os.access("/dev/shm", os.W_OK) else "/tmp"


def write_file(path, arr, n, copies):
    """`copies` copies of the block arr[:n] as one file"""
    with open(path, "wb") as f:
        for _ in range(copies):
            arr[:n].tofile(f)
    return n * copies


def uri_of(path, shape):
    if shape["fmt"] == "csv":
        return "%s?format=csv&label_column=%d" % (path, shape["lc"])
    return path


def algorithmic_bytes(n, blk, ib):
    """every input byte read once, every output element written once"""
    return (n + 8 * (blk.rows + 1) + 4 * blk.rows + 4 * blk.n_weight + 8 * blk.n_qid +
            blk.nnz * ib + blk.n_field * ib + 4 * blk.n_value)


# ------------------------------------------------------------------------------------------------
def reference_rate(ref, shape, path, nthread, seconds, part=0, nparts=1, passes=0):
    """The unmodified reference: Parser<I>::Create(uri, part, nparts, type) drained, sizes only, in a
    process of its own with OMP_NUM_THREADS = nthread (oracle/ref_time.py says why)."""
    if ref is None:
        return None
    env = dict(os.environ, OMP_NUM_THREADS=str(nthread))
    cmd = [sys.executable, os.path.join(ROOT, "oracle", "ref_time.py"), uri_of(path, shape), str(part),
           str(nparts), shape["fmt"], str(shape["ib"]), str(seconds)] + ([str(passes)] if passes else [])
    r = subprocess.run(cmd, env=env, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL, text=True)
    try:
        got = json.loads(r.stdout.strip().splitlines()[-1])
    except Exception:
        return None
    return None if "error" in got else got


def run_reference_arm(args, out):
    """--impl reference: the reference's own multithreaded CPU parser on this box's host cores, on
    the headline shape (and, briefly, on the other four)."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from dmlc_core_b200 import datagen
    from oracle.pyoracle import Reference
    ref = Reference.load()
    if ref is None:
        out.write(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libdmlc_ref.so missing"}) + "\n")
        out.flush()
        return
    nthread = os.cpu_count() or 1
    shape = SHAPES[args.shape]
    sample_mb = 1000
    kw = {"ncols": 128} if shape["kind"] == "csv" else {}
    data, rows = datagen.generate(shape["kind"], 1, sample_mb * 1000 * 1000, **kw)
    path = os.path.join(shm_dir(), "b200_ref_sample.txt")
    data.tofile(path)
    uri = uri_of(path, shape)
    nbytes = data.size
    del data
    # one process of its own per leg, OMP_NUM_THREADS = all host threads: the reference sizes its
    # OpenMP team from that variable alone, and launchers (torchrun) export OMP_NUM_THREADS=1
    got = reference_rate(ref, shape, path, nthread, 0.0, passes=args.steps)   # behind one untimed pass
    os.remove(path)
    if got is None:
        out.write(json.dumps({"impl": "reference", "unavailable": "the reference failed on the sample"}) + "\n")
        out.flush()
        return
    ms = got["seconds_per_pass"] * 1e3
    value = got["GBps"]
    r = {"rows": got["rows"]}
    t = got["seconds_per_pass"] * args.steps
    shapes = []
    if not args.no_shapes:
        for i, sh in enumerate(SHAPES):
            if i == args.shape:
                shapes.append({"config": sh["key"], "reference_GBps": value, "threads": nthread})
                continue
            kw = {"ncols": 128} if sh["kind"] == "csv" else {}
            d, _ = datagen.generate(sh["kind"], 1, 300 * 1000 * 1000, **kw)
            p = os.path.join(shm_dir(), "b200_ref_shape.txt")
            d.tofile(p)
            del d
            got = reference_rate(ref, sh, p, nthread, 2.0)
            os.remove(p)
            shapes.append({"config": sh["key"], "reference_GBps": got["GBps"] if got else None,
                           "threads": nthread, "sample_MB": 300})
    sample = ("%d MB of the same seeded %s text (first block of the workload) read from a file in %s "
              "through Parser::Create, 1 pass per step" % (sample_mb, shape["fmt"], shm_dir()))
    line = {
        "impl": "reference", "metric": "parsed input GB/s", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": max(1, args.warmup), "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64+f32",
        "data": "synthetic", "rows_per_s": r["rows"] / (t / args.steps),
        "config": {"workload": "%s: %s; reference CPU parser (all %d host threads) on a bounded %d MB sample"
                               % (shape["key"], shape["what"], nthread, sample_mb),
                   "source": "file in %s (page cache), dmlc::Parser<IndexType>::Create -> Next()/Value()" % shm_dir()},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": nthread, "kind": "reference", "sample": sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    if shapes:
        line["shapes"] = shapes
    out.write(json.dumps(line) + "\n")
    out.flush()


# ------------------------------------------------------------------------------------------------
class DeviceBench(object):
    """Device-resident measurement of one shape: `copies` windows of one block per step, the
    windows alternating between `streams` handles (one chunk per handle in flight, as the ABI asks)."""

    def __init__(self, torch, shape, dev_ptrs, n, local, streams, range_bytes):
        from dmlc_core_b200 import Parser
        self.torch = torch
        self.ptrs = dev_ptrs
        self.n = n
        self.main = torch.cuda.current_stream()
        self.lanes = [(Parser(shape["fmt"], index_bytes=shape["ib"], label_column=shape["lc"], device=local,
                              range_bytes=range_bytes, stream=self.main.cuda_stream), self.main)]
        for _ in range(max(1, streams) - 1):
            side = torch.cuda.Stream()
            self.lanes.append((Parser(shape["fmt"], index_bytes=shape["ib"], label_column=shape["lc"],
                                      device=local, range_bytes=range_bytes, stream=side.cuda_stream), side))

    def step(self, lanes=None):
        torch = self.torch
        lanes = lanes or self.lanes
        fork = torch.cuda.Event()
        fork.record(self.main)
        for (_, st) in lanes[1:]:
            st.wait_event(fork)
        pending = [False] * len(lanes)
        blk = None
        for j, ptr in enumerate(self.ptrs):
            k = j % len(lanes)
            if pending[k]:
                blk = lanes[k][0].parse_device_finish()     # every window is finished and checked
            lanes[k][0].parse_device_async(ptr, self.n)
            pending[k] = True
        for k in range(len(lanes) - 1, -1, -1):              # lane 0 (the main stream) last
            if pending[k]:
                blk = lanes[k][0].parse_device_finish()
            if k != 0:
                done = torch.cuda.Event()
                done.record(lanes[k][1])
                self.main.wait_event(done)
        return blk

    def launches(self):
        return sum(q.launch_count() for (q, _) in self.lanes)

    def kernel_ms(self, reps=2):
        p = self.lanes[0][0]
        p.set_profiling(True)
        for _ in range(reps):
            self.step(self.lanes[:1])        # one stream: per-kernel times must not overlap other launches
        self.torch.cuda.synchronize()
        kt = [p.kernel_time(k) for k in range(3)]
        p.set_profiling(False)
        return [kt[k][0] / max(kt[k][1], 1) if kt[k][1] else 0.0 for k in range(3)]

    def close(self):
        for (q, _) in self.lanes:
            q.close()


def e2e_cabi(torch, shape, host, n, copies, args, local, out_device=False, steps=3):
    """pinned host text -> CSR through b200p_submit / b200p_next, `copies` passes over the block per step"""
    from dmlc_core_b200 import Parser
    harr = host.numpy()
    chunks = cut_chunks(harr, n, args.chunk_mb << 20)
    pe = Parser(shape["fmt"], index_bytes=shape["ib"], label_column=shape["lc"], device=local,
                range_bytes=args.range_bytes, chunk_bytes=args.chunk_mb << 20, depth=args.depth,
                out_device=out_device)
    base_ptr = harr.ctypes.data
    if args.host_threads >= 0:
        pe.set_host_threads(args.host_threads)
    handed = [0]

    def take():
        r = pe.next_raw()
        handed[0] += (8 * (r.rows + 1) + 4 * r.rows + 4 * r.n_weight + 8 * r.n_qid +
                      shape["ib"] * (r.nnz + r.n_field) + 4 * r.n_value)
        return r.rows

    def one_step():
        got_rows = 0
        for _ in range(copies):
            for (a, b) in chunks:
                while pe.pending() >= args.depth:
                    got_rows += take()
                pe.submit_in_place(base_ptr + a, b - a)
        while pe.pending():
            got_rows += take()
        return got_rows

    return pe, one_step, handed


def link_probe(torch, dist, world, host, dev, nbytes_block=64 << 20, reps=8):
    """Copy-only probe of the host link, every rank at the same time (barrier before each leg):
    H2D only, D2H only, both at once, 64 MiB blocks like the pipeline.  Per-rank rates and, summed
    over the ranks, what the whole box moves -- the ceiling the end-to-end numbers sit under."""
    nb = min(nbytes_block, host.numel())
    d_tmp = torch.empty(nb, dtype=torch.uint8, device="cuda")
    h_out = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
    s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

    def leg(up, down):
        torch.cuda.synchronize()
        if world > 1:
            dist.barrier()
        t0 = time.perf_counter()
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s_up):
                    d_tmp.copy_(host[:nb], non_blocking=True)
            if down:
                with torch.cuda.stream(s_dn):
                    h_out.copy_(dev[:nb], non_blocking=True)
        torch.cuda.synchronize()
        dt = time.perf_counter() - t0
        if world > 1:
            t = torch.tensor([dt], dtype=torch.float64, device="cuda")
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            dt_max = float(t.item())
        else:
            dt_max = dt
        return reps * nb / dt / GB, world * reps * nb / dt_max / GB

    leg(True, True)
    up, up_box = leg(True, False)
    down, down_box = leg(False, True)
    both, both_box = leg(True, True)
    return ({"h2d_only_GBps": up, "d2h_only_GBps": down, "h2d_with_concurrent_d2h_GBps": both,
             "how": "pinned cudaMemcpyAsync of %d MiB blocks, %d reps, this rank" % (nb >> 20, reps)},
            {"h2d_only_GBps": up_box, "d2h_only_GBps": down_box, "each_direction_when_both_GBps": both_box,
             "ranks": world, "how": "the same legs started behind a barrier on every rank; bytes of all ranks / slowest rank's time"})


def measure_shape(torch, shape, args, local, peak, ref, nthread):
    """One BASELINE config at a bounded size: device-resident per-kernel times and rate, end to end
    through the C ABI, and the reference CPU parser on a sample of the same text."""
    host, n, rows = make_block(shape, args.shape_mb, pinned=True)
    dev = host[:n].cuda()
    db = DeviceBench(torch, shape, [dev.data_ptr()], n, local, 1, args.range_bytes)
    blk = None
    for _ in range(3):
        blk = db.step()
    assert int(blk.rows) == rows, (shape["key"], int(blk.rows), rows)
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    reps = 5
    ev0.record(db.main)
    for _ in range(reps):
        db.step()
    ev1.record(db.main)
    torch.cuda.synchronize()
    ms = ev0.elapsed_time(ev1) / reps
    kms = db.kernel_ms()
    b_alg = algorithmic_bytes(n, blk, shape["ib"])
    total_k = sum(kms)
    dominant = max(kms)
    out = {"config": shape["key"], "what": shape["what"], "named_size_GB": shape["gb"],
           "measured_on_bytes": n, "rows": int(blk.rows), "nnz": int(blk.nnz),
           "value": n / (ms * 1e-3) / GB, "unit": "GB/s", "ms_per_pass": ms,
           "kernel_ms": {"collect_or_count": kms[0], "scan": kms[1], "convert_or_parse_or_emit": kms[2]},
           "path": {0: "two-pass walk", 1: "one launch (ParseTiles)", 2: "collect / scan / convert"}[db.lanes[0][0].last_path()],
           "roofline": {"bound": "hbm", "algorithmic_bytes": int(b_alg),
                        "dominant_kernel_frac": b_alg / (dominant * 1e-3) / GB / peak if dominant else None,
                        "all_kernels_frac": b_alg / (total_k * 1e-3) / GB / peak if total_k else None,
                        "peak": peak}}
    db.close()
    del dev
    if not args.no_e2e:
        pe, one_step, handed = e2e_cabi(torch, shape, host, n, 1, args, local)
        assert one_step() == rows
        one_step()
        up0, down0 = pe.copied_bytes()
        t0 = time.perf_counter()
        for _ in range(3):
            one_step()
        dt = (time.perf_counter() - t0) / 3
        up1, down1 = pe.copied_bytes()
        pe.close()
        out["e2e"] = {"value": n / dt / GB, "unit": "GB/s", "h2d_bytes_per_step": (up1 - up0) // 3,
                      "d2h_bytes_per_step": (down1 - down0) // 3}
    if ref is not None:
        harr = host.numpy()
        sample = chunks_end_at(harr, min(n, 300 * 1000 * 1000))
        path = os.path.join(shm_dir(), "b200_bench_shape.txt")
        harr[:sample].tofile(path)
        got = reference_rate(ref, shape, path, nthread, 2.0)
        os.remove(path)
        if got is not None:
            out["reference_cpu"] = {"value": got["GBps"], "unit": "GB/s", "cores": nthread,
                                    "sample": "first %d MB of the same text, from a file, Parser::Create" % (sample // 1000000)}
    del host
    return out


def claim_stdout():
    """The contract is ONE JSON line on stdout.  Libraries print there too (NCCL's version banner at
    any NCCL_DEBUG level from VERSION up, whatever NCCL_DEBUG_FILE says), so everything that is not
    the line goes to stderr: file descriptor 1 is pointed at stderr and the line is written to a
    private copy of the original stdout."""
    sys.stdout.flush()
    keep = os.dup(1)
    os.dup2(2, 1)
    return os.fdopen(keep, "w")


def main():
    args = parse_args()
    out = claim_stdout()
    # launchers export OMP_NUM_THREADS=1; the text generator (OpenMP, csrc/datagen.c) may use this
    # rank's share of the host cores
    ranks_here = max(1, int(os.environ.get("LOCAL_WORLD_SIZE", "1")))
    os.environ["OMP_NUM_THREADS"] = str(max(1, (os.cpu_count() or 1) // ranks_here))
    if args.impl == "reference":
        run_reference_arm(args, out)
        return
        run_reference_arm(args)
        return
    import torch
    import torch.distributed as dist
    from dmlc_core_b200 import hostapi, build

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (there is no CPU path)")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout must carry exactly one JSON line; NCCL's own log goes to stderr, at whatever level
        # the launcher asked for (NCCL_DEBUG is left alone)
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def reduce(x, op):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=op)
        return float(t.item())

    def max_over_ranks(x):
        return reduce(x, dist.ReduceOp.MAX) if world > 1 else x

    def sum_over_ranks(x):
        return reduce(x, dist.ReduceOp.SUM) if world > 1 else x

    shape = SHAPES[args.shape]
    ib = shape["ib"]
    peak, peak_src = measured_peak()
    nthread = os.cpu_count() or 1
    os.environ["DMLC_B200_DEVICE"] = str(local)      # the C++ layer of this rank parses on this rank's GPU

    # ---------------- workload --------------------------------------------------------------
    t_setup = time.time()
    host, n, rows_block = make_block(shape, args.block_mb, pinned=True)
    want_gb = args.gb if args.gb > 0 else min(shape["gb"], 20.0)
    copies = max(1, int(round(want_gb * GB / n)))
    dev_block = host[:n].cuda(non_blocking=False)
    dev = torch.empty(copies * n, dtype=torch.uint8, device="cuda")
    for j in range(copies):
        dev[j * n:(j + 1) * n].copy_(dev_block)
    del dev_block
    text_bytes = copies * n
    harr = host.numpy()
    setup_s = time.time() - t_setup

    db = DeviceBench(torch, shape, [dev.data_ptr() + j * n for j in range(copies)], n, local,
                     args.streams, args.range_bytes)

    # ---------------- device-resident: value --------------------------------------------------
    # clocks are sampled from the warm-up on: the timed region itself is shorter than a few of
    # nvidia-smi's polling periods, and the warm-up runs the very same kernels back to back
    sampler = ClockSampler(local)
    sampler.start()
    blk = None
    for _ in range(max(args.warmup, 3)):
        blk = db.step()
    rows, nnz = int(blk.rows), int(blk.nnz)
    assert rows == rows_block, (rows, rows_block)
    launches0 = db.launches()
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(db.main)
    for _ in range(args.steps):
        db.step()
    ev1.record(db.main)
    launches = db.launches() - launches0     # kernels enqueued inside the timed region
    barrier()
    t_extra = time.time()
    while len(sampler.lines) < 4 and time.time() - t_extra < 3.0:
        db.step()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    clocks["window"] = "warm-up + timed steps + identical untimed steps until >= 4 samples"
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    value = world * text_bytes / (dev_ms * 1e-3) / GB

    # ---------------- roofline: the kernels of one chunk against the HBM copy peak -----------------
    kms = db.kernel_ms()
    b_alg = algorithmic_bytes(n, blk, ib)
    dominant_ms = max(kms)
    total_ms = sum(kms)
    fmt_id = {"libsvm": 0, "libfm": 1, "csv": 2}[shape["fmt"]]
    idx_t = "unsigned long" if ib == 8 else "unsigned int"
    path = db.lanes[0][0].last_path()
    names = {2: ["CollectRanges<%d, %s" % (fmt_id, idx_t), "ScanTiles+ScanFinish", "ConvertRanges<%d, %s" % (fmt_id, idx_t)],
             1: ["", "", "ParseTiles<%d, %s" % (fmt_id, idx_t)],
             0: ["WalkText<%d, %s, 0" % (fmt_id, idx_t), "ScanTiles+ScanFinish", "WalkText<%d, %s, 1" % (fmt_id, idx_t)]}[path]
    dominant = names[kms.index(dominant_ms)]
    # The path's algorithmic bytes (text read once, every array element written once) are
    # the work of ALL its launches on a chunk, so the headline fraction divides them by the summed
    # kernel time; the dominant launch is shown beside it with the bytes it moves itself.
    traffic = [ncu_traffic(k) for k in names if "<" in k]       # the two scan launches move kilobytes
    scaled = [t["bytes_per_launch"] * (n / t["captured_on_bytes"]) if t and t.get("captured_on_bytes") else
              (t["bytes_per_launch"] if t else None) for t in traffic]
    achieved = b_alg / (total_ms * 1e-3) / GB
    roofline = {"bound": "hbm",
                "kernel": " + ".join(k + ", ...>" if "<" in k else k for k in names if k),
                "achieved": achieved, "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": sum(scaled) if scaled and all(x is not None for x in scaled) else None,
                "traffic_source": traffic,
                "peak_source": peak_src, "algorithmic_bytes_per_launch": int(b_alg),
                "launch_ms": total_ms,
                "path": {0: "two-pass walk", 1: "one launch", 2: "collect / scan / convert"}[path],
                "kernels": [{"name": k, "ms": ms} for k, ms in zip(names, kms) if k],
                "dominant": {"kernel": dominant + ", ...>", "ms": dominant_ms,
                             "share_of_kernel_time": dominant_ms / total_ms,
                             "path_bytes_over_its_time": {"achieved": b_alg / (dominant_ms * 1e-3) / GB,
                                                          "frac": b_alg / (dominant_ms * 1e-3) / GB / peak}},
                "whole_step": {"ms_per_window": dev_ms / copies,
                               "frac": b_alg / (dev_ms / copies * 1e-3) / GB / peak,
                               "note": "windows alternate between %d streams, so launches of neighbouring windows overlap" % len(db.lanes)}}

    # ---------------- the link the end-to-end numbers ride on ---------------------------------------
    link = link_box = None
    if not args.no_e2e:
        link, link_box = link_probe(torch, dist, world, host, dev)

    # ---------------- end to end through the C ABI: pinned host text -> pinned host CSR -----------
    e2e = e2e_dev = None
    if not args.no_e2e:
        e2e_steps = max(1, min(args.steps, 3))
        e2e_copies = max(1, min(copies, 10))        # 10 GB per step is plenty for a stable rate
        for out_device in (False, True):
            pe, one_step, handed = e2e_cabi(torch, shape, host, n, e2e_copies, args, local, out_device)
            assert one_step() == rows * e2e_copies
            one_step()
            barrier()
            up0, down0 = pe.copied_bytes()
            handed[0] = 0
            t0 = time.perf_counter()
            for _ in range(e2e_steps):
                one_step()
            torch.cuda.synchronize()
            e2e_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
            up1, down1 = pe.copied_bytes()
            barrier()
            pe.close()
            res = {"value": world * e2e_copies * n / e2e_s / GB, "unit": "GB/s",
                   "h2d_bytes_per_step": (up1 - up0) // e2e_steps,
                   "d2h_bytes_per_step": (down1 - down0) // e2e_steps,
                   "result_bytes_per_step": handed[0] // e2e_steps,
                   "text_bytes_per_step_per_gpu": e2e_copies * n,
                   "ms_per_step": e2e_s * 1e3, "chunk_bytes": args.chunk_mb << 20, "depth": args.depth}
            if out_device:
                res["api"] = "b200p_submit / b200p_next with out_residency = device (pinned host text in, CSR stays in HBM)"
                res["frac_of_box_h2d"] = res["value"] / link_box["h2d_only_GBps"]
                e2e_dev = res
            else:
                res["api"] = "b200p_submit / b200p_next (pinned host text in, pinned host CSR out)"
                res["source"] = "text already in pinned host memory (no file read inside the timed region)"
                res["host_threads"] = args.host_threads if args.host_threads >= 0 else "library default (<= 4)"
                res["link"] = link
                res["link_box"] = link_box
                res["frac_of_h2d_link_under_duplex"] = (e2e_copies * n / e2e_s / GB) / link["h2d_with_concurrent_d2h_GBps"]
                res["frac_of_box_duplex"] = res["value"] / link_box["each_direction_when_both_GBps"]
                e2e = res

    db.close()
    del dev
    torch.cuda.empty_cache()

    # ---------------- the drop-in API on a file: Parser<I>::Create(uri, rank, N, type) -------------
    e2e_create = parity = cpu = cpu_sharded = None
    ref = None
    if not args.no_cpu_baseline:
        from oracle.pyoracle import Reference
        ref = Reference.load()
    if not args.no_e2e:
        build.build_host()
        hostapi.load()
        file_copies = max(1, int(round(args.file_gb * GB * world / n)))
        path = os.path.join(shm_dir(), "b200_bench_shared.txt")
        if rank == 0:
            try:
                write_file(path, harr, n, file_copies)
            except OSError:
                file_copies = 1
                write_file(path, harr, n, file_copies)
        if world > 1:
            t = torch.tensor([file_copies], dtype=torch.int64, device="cuda")
            dist.broadcast(t, 0)
            file_copies = int(t.item())
        barrier()
        file_bytes = n * file_copies
        uri = uri_of(path, shape)
        passes = 3
        hostapi.time_parser(uri, rank, world, shape["fmt"], ib, passes=1)       # warm: allocations, page cache
        barrier()
        t0 = time.perf_counter()
        got = hostapi.time_parser(uri, rank, world, shape["fmt"], ib, passes=passes)
        wall = max_over_ranks(time.perf_counter() - t0)
        secs = max_over_ranks(got["seconds"])
        rows_all = sum_over_ranks(got["rows"])
        bytes_all = sum_over_ranks(got["bytes_read"])
        assert int(rows_all) == rows_block * file_copies * passes, (rows_all, rows_block, file_copies)
        assert int(bytes_all) == file_bytes * passes, (bytes_all, file_bytes)
        e2e_create = {"value": file_bytes * passes / secs / GB, "unit": "GB/s",
                      "api": "dmlc::Parser<%s>::Create(uri, part_index=rank, num_parts=%d, \"%s\") -> Next()/Value(), "
                             "BeforeFirst between the %d passes (libdmlc_b200_host.so)"
                             % ("uint64_t" if ib == 8 else "uint32_t", world, shape["fmt"], passes),
                      "source": "one file of %d bytes in %s (page cache), shared by all ranks" % (file_bytes, shm_dir()),
                      "parts": world, "rows_all_ranks_per_pass": int(rows_all) // passes,
                      "bytes_read_all_ranks_per_pass": int(bytes_all) // passes,
                      "seconds_incl_create": wall, "sum_of_rows_and_bytes_matches_the_file": True}
        # the CPU side of the same split: N processes x (cores / N) threads, part_index = rank
        if ref is not None and world > 1:
            barrier()
            t0 = time.perf_counter()
            r = reference_rate(ref, shape, path, max(1, nthread // world), 0.0, rank, world, passes=1)
            dt = max_over_ranks(r["seconds_per_pass"] if r else 0.0)
            assert r is not None and int(sum_over_ranks(r["rows"])) == rows_block * file_copies
            cpu_sharded = {"value": file_bytes / dt / GB, "unit": "GB/s", "processes": world,
                           "wall_incl_process_start_s": max_over_ranks(time.perf_counter() - t0),
                           "threads_per_process": max(1, nthread // world), "kind": "reference",
                           "sample": "the same shared file, one pass, part_index = rank of %d" % world}
        # the timed text is checked against the reference: hashes of the seven arrays over a sample
        if rank == 0 and ref is not None:
            sample_bytes = chunks_end_at(harr, min(n, 200 * 1000 * 1000))
            spath = os.path.join(shm_dir(), "b200_bench_parity.txt")
            harr[:sample_bytes].tofile(spath)
            suri = uri_of(spath, shape)
            mine = hostapi.time_parser(suri, 0, 1, shape["fmt"], ib, passes=1, want_hash=True)
            theirs, ref_rows, ref_nnz = ref.hash_uri(suri, 0, 1, shape["fmt"], ib, nthread=nthread)
            os.remove(spath)
            parity = {"equal": mine["hash"] == theirs and mine["rows"] == ref_rows and mine["nnz"] == ref_nnz,
                      "arrays": ["row lengths", "label", "weight", "qid", "field", "index", "value"],
                      "sample": "first %d bytes of the timed block, both sides through Parser::Create on one file; "
                                "FNV-1a 64 per array" % sample_bytes,
                      "rows": ref_rows, "nnz": ref_nnz}
        barrier()
        if rank == 0:
            os.remove(path)

    # ---------------- CPU baseline (rank 0, N == 1) -------------------------------------------
    if rank == 0 and world == 1 and ref is not None:
        sample_bytes = chunks_end_at(harr, min(n, 500 * 1000 * 1000))
        path = os.path.join(shm_dir(), "b200_bench_sample.txt")
        harr[:sample_bytes].tofile(path)
        res = reference_rate(ref, shape, path, nthread, args.cpu_seconds)
        one = reference_rate(ref, shape, path, 1, min(4.0, args.cpu_seconds)) if res is not None and nthread > 1 else None
        os.remove(path)
        if res is not None:
            cpu = {"value": res["GBps"], "unit": "GB/s", "cores": nthread, "kind": "reference",
                   "sample": "first %d MB of the workload from a file in %s, %.1f s of passes, OMP threads=%d"
                             % (sample_bytes // 1000000, shm_dir(), args.cpu_seconds, nthread),
                   "rows_per_s": res["rows_per_s"]}
            if one is not None:   # OMP_NUM_THREADS in {1, all cores}
                cpu["single_thread_GBps"] = one["GBps"]

    # ---------------- all five BASELINE configs (N == 1) -----------------------------------------
    shapes = None
    if world == 1 and not args.no_shapes:
        del host
        torch.cuda.empty_cache()
        shapes = []
        for i, sh in enumerate(SHAPES):
            shapes.append(measure_shape(torch, sh, args, local, peak, ref, nthread))

    if rank == 0:
        line = {
            "metric": "parsed input GB/s", "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": dev_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "u64+f32" if ib == 8 else "u32+f32",
            "data": "synthetic (seeded generator; one unique %d MB block tiled %dx per GPU)"
                    % (n // 1000000, copies),
            "config": {"workload": "%s: %s, 1xB200 per rank" % (shape["key"], shape["what"]),
                       "text_bytes_per_gpu": text_bytes, "rows_per_gpu": rows * copies,
                       "nnz_per_gpu": nnz * copies, "window_bytes": n,
                       "l2": "inputs (%d GB) far larger than the 126 MB L2; no flush needed" % (text_bytes // 1000000000),
                       "parallelism": "shard-per-GPU, no collective",
                       "streams": len(db.lanes),
                       "source": "device-resident `value`: text in HBM; `e2e`: text in pinned host memory; "
                                 "`e2e_parser_create`: a file in %s through Parser::Create" % shm_dir()},
            "rows_per_s": world * rows * copies / (dev_ms * 1e-3),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "setup_s": setup_s,
        }
        if e2e is not None:
            line["e2e"] = e2e
        if e2e_dev is not None:
            line["e2e_device_out"] = e2e_dev
        if e2e_create is not None:
            line["e2e_parser_create"] = e2e_create
        if parity is not None:
            line["parity_in_bench"] = parity["equal"]
            line["parity_detail"] = parity
        if cpu is not None:
            line["cpu_baseline"] = cpu
        if cpu_sharded is not None:
            line["cpu_sharded"] = cpu_sharded
        if shapes is not None:
            line["shapes"] = shapes
        out.write(json.dumps(line) + "\n")
        out.flush()
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
