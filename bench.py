#!/usr/bin/env python
"""bench.py -- parsed input GB/s of the B200 text -> CSR path (BASELINE.json metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl reference]

Workload (N=1): BASELINE.json configs[1] -- dense CSV, 10 GB synthetic, 128 float columns + a
label column, Parser<uint32_t>, one B200.  One step = one pass of the hot path over the whole
10 GB.  With N > 1 (torchrun, one rank per GPU) every rank parses its own 10 GB shard with no
collective on the data path ("weak" scaling, like InputSplit part_index / num_parts sharding).

Printed JSON (rank 0):
  value     whole-job GB/s of input text, text already resident in HBM (CUDA events)
  e2e       same metric through b200p_submit / b200p_next with the text in pinned HOST memory:
            H2D of every chunk and D2H of every CSR array inside the timed region
  roofline  the dominant kernel (pass 2, WalkText<EMIT>) against the measured HBM copy peak
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
CSV_COLS = 128
LABEL_COLUMN = 0


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="b200")
    ap.add_argument("--gb", type=float, default=10.0, help="text bytes per GPU per step (GB)")
    ap.add_argument("--block-mb", type=int, default=1000, help="unique generated block (MB)")
    ap.add_argument("--chunk-mb", type=int, default=64, help="e2e chunk size (MiB)")
    ap.add_argument("--range-bytes", type=int, default=0)
    ap.add_argument("--depth", type=int, default=3, help="e2e chunks in flight")
    ap.add_argument("--streams", type=int, default=2,
                    help="device-resident measurement: handles/streams the windows of a step alternate between")
    ap.add_argument("--host-threads", type=int, default=-1,
                    help="e2e: b200p_set_host_threads (-1 = library default, 0 = copy index[] back)")
    ap.add_argument("--cpu-seconds", type=float, default=12.0)
    ap.add_argument("--no-cpu-baseline", action="store_true")
    ap.add_argument("--no-e2e", action="store_true")
    return ap.parse_args()


def ncu_traffic(kernel_substr):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the
    committed `ncu --set full` capture of this same workload (profiles/current_ncu.metrics.json,
    written by profiles/summarize_ncu.py). bench.py itself never runs under a profiler."""
    try:
        with open(os.path.join(ROOT, "profiles", "current_ncu.metrics.json")) as f:
            for k in json.load(f):
                if kernel_substr in k.get("kernel", ""):
                    return float(k["dram_bytes_per_launch"])
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


def make_block(block_mb, pinned):
    """One unique block of dense CSV (whole lines), padded with '\\n' to a multiple of 16 bytes
    (blank lines are skipped by the reference, csv_parser.h:72,98) so tiled copies stay aligned."""
    import torch
    from dmlc_core_b200 import datagen
    target = block_mb * 1000 * 1000
    cap = target + (8 << 20)
    host = torch.empty(cap, dtype=torch.uint8, pin_memory=pinned)
    arr = host.numpy()
    data, rows = datagen.generate("csv", 1, target, ncols=CSV_COLS, out=arr)
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


def time_reference(sample_path, nthread, seconds, steps=None):
    """The unmodified reference (oracle/_ref): Parser<uint32_t>::Create(uri, 0, 1, "csv") drain."""
    from oracle.pyoracle import Reference
    ref = Reference.load()
    if ref is None:
        return None
    uri = "%s?format=csv&label_column=%d" % (sample_path, LABEL_COLUMN)
    nbytes = os.path.getsize(sample_path)
    first = ref.time_uri(uri, 0, 1, "csv", 4, nthread=nthread, passes=1)
    if first["seconds"] <= 0:
        return None
    if steps is None:
        passes = max(1, int(seconds / first["seconds"]))
        res = ref.time_uri(uri, 0, 1, "csv", 4, nthread=nthread, passes=passes)
        per = res["seconds"] / passes
    else:
        per = None
    return {"ref": ref, "uri": uri, "bytes": nbytes, "first": first, "per_pass": per,
            "rows": first["rows"]}


def run_reference_arm(args):
    """--impl reference: the reference's own multithreaded CPU parser on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    from dmlc_core_b200 import datagen
    from oracle.pyoracle import Reference
    ref = Reference.load()
    kind = "reference"
    if ref is None:
        print(json.dumps({"impl": "reference", "unavailable": "oracle/_ref/libdmlc_ref.so missing"}))
        return
    nthread = os.cpu_count() or 1
    sample_mb = 1000
    data, rows = datagen.generate("csv", 1, sample_mb * 1000 * 1000, ncols=CSV_COLS)
    path = "/dev/shm/b200_ref_sample.csv" if os.path.isdir("/dev/shm") else "/tmp/b200_ref_sample.csv"
    data.tofile(path)
    uri = "%s?format=csv&label_column=%d" % (path, LABEL_COLUMN)
    nbytes = data.size
    del data
    for _ in range(args.warmup):
        ref.time_uri(uri, 0, 1, "csv", 4, nthread=nthread, passes=1)
    t = 0.0
    for _ in range(args.steps):
        r = ref.time_uri(uri, 0, 1, "csv", 4, nthread=nthread, passes=1)
        t += r["seconds"]
    os.remove(path)
    ms = t / args.steps * 1e3
    value = nbytes / (t / args.steps) / GB
    sample = "%d MB of the same seeded CSV (first block of the workload), 1 pass per step" % sample_mb
    line = {
        "impl": "reference", "metric": "parsed input GB/s", "value": value, "unit": "GB/s",
        "n_gpus": args.gpus, "steps": args.steps, "warmup": args.warmup, "ms_per_step": ms,
        "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
        "data": "synthetic", "rows_per_s": r["rows"] / (t / args.steps),
        "config": {"workload": "dense CSV parse, 128 float columns + label_column=0, Parser<uint32_t>; "
                               "reference CPU parser on a bounded %d MB sample" % sample_mb},
        "cpu_baseline": {"value": value, "unit": "GB/s", "cores": nthread, "kind": kind, "sample": sample},
        "e2e": {"value": value, "unit": "GB/s", "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
        "gpu_launches": 0,
    }
    print(json.dumps(line))


def main():
    args = parse_args()
    if args.impl == "reference":
        run_reference_arm(args)
        return
    import torch
    import torch.distributed as dist
    from dmlc_core_b200 import Parser

    world = int(os.environ.get("WORLD_SIZE", "1"))
    rank = int(os.environ.get("RANK", "0"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py needs a GPU (there is no CPU path)")
    torch.cuda.set_device(local)
    if world > 1:
        # stdout must carry exactly one JSON line: NCCL prints its version banner there at any
        # debug level from VERSION up (WARN included), so switch its logging off and point
        # whatever is left at stderr
        os.environ.pop("NCCL_DEBUG", None)
        os.environ["NCCL_DEBUG_FILE"] = "/dev/stderr"
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x):
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    # ---------------- workload --------------------------------------------------------------
    t_setup = time.time()
    host, n, rows_block = make_block(args.block_mb, pinned=True)
    copies = max(1, int(round(args.gb * GB / n)))
    dev_block = host[:n].cuda(non_blocking=False)
    dev = torch.empty(copies * n, dtype=torch.uint8, device="cuda")
    for j in range(copies):
        dev[j * n:(j + 1) * n].copy_(dev_block)
    del dev_block
    text_bytes = copies * n
    harr = host.numpy()
    chunks = cut_chunks(harr, n, args.chunk_mb << 20)
    setup_s = time.time() - t_setup

    stream = torch.cuda.current_stream()
    p = Parser("csv", index_bytes=4, label_column=LABEL_COLUMN, device=local,
               range_bytes=args.range_bytes, stream=stream.cuda_stream)
    # The windows of a step are independent chunks.  With --streams 2 (default) they alternate
    # between two handles on two streams, so that pass 1 of one window (bound by HBM reads) runs
    # beside pass 2 of another (bound by instruction issue).
    # All timing stays on the main stream: the side streams fork from it and join it by events.
    lanes = [(p, stream)]
    for _ in range(max(1, args.streams) - 1):
        side = torch.cuda.Stream()
        lanes.append((Parser("csv", index_bytes=4, label_column=LABEL_COLUMN, device=local,
                             range_bytes=args.range_bytes, stream=side.cuda_stream), side))

    def device_step(lanes=lanes):
        fork = torch.cuda.Event()
        fork.record(stream)
        for (_, st) in lanes[1:]:
            st.wait_event(fork)
        used = set()
        for j in range(copies):
            k = j % len(lanes)
            lanes[k][0].parse_device_async(dev.data_ptr() + j * n, n)
            used.add(k)
        blk = None
        for k in sorted(used, reverse=True):     # lane 0 (the main stream) last
            blk = lanes[k][0].parse_device_finish()
            if k != 0:
                done = torch.cuda.Event()
                done.record(lanes[k][1])
                stream.wait_event(done)
        return blk

    # ---------------- device-resident: value --------------------------------------------------
    # clocks are sampled from the warm-up on: the timed region itself (tens of ms) is shorter than
    # nvidia-smi's polling period, and the warm-up runs the very same kernels back to back
    sampler = ClockSampler(local)
    sampler.start()
    for _ in range(max(args.warmup, 3)):
        blk = device_step()
    rows = int(blk.rows)
    nnz = int(blk.nnz)
    assert rows == rows_block and nnz == rows * CSV_COLS, (rows, rows_block, nnz)
    launches0 = sum(q.launch_count() for (q, _) in lanes)
    barrier()
    ev0, ev1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    ev0.record(stream)
    for _ in range(args.steps):
        device_step()
    ev1.record(stream)
    launches = sum(q.launch_count() for (q, _) in lanes) - launches0   # kernels enqueued inside the timed region
    barrier()
    # the timed region lasts tens of ms; keep the identical load running (untimed) until
    # nvidia-smi has delivered a few samples
    t_extra = time.time()
    while len(sampler.lines) < 4 and time.time() - t_extra < 3.0:
        device_step()
    torch.cuda.synchronize()
    clocks = sampler.stop()
    clocks["window"] = "warm-up + timed steps + identical untimed steps until >= 4 samples"
    dev_ms = max_over_ranks(ev0.elapsed_time(ev1)) / args.steps
    value = world * text_bytes / (dev_ms * 1e-3) / GB

    # ---------------- roofline of the dominant kernel (pass 2) ---------------------------------
    p.set_profiling(True)
    for _ in range(2):
        device_step(lanes[:1])      # one stream: per-kernel times must not overlap other launches
    torch.cuda.synchronize()
    kt = [p.kernel_time(k) for k in range(3)]
    p.set_profiling(False)
    emit_ms = kt[2][0] / max(kt[2][1], 1)
    count_ms = kt[0][0] / max(kt[0][1], 1)
    scan_ms = kt[1][0] / max(kt[1][1], 1)
    # algorithmic bytes of one launch (one ~1 GB window): SURVEY 8(d)
    b_alg = n + 8 * (rows + 1) + 4 * rows + nnz * (4 + 4)
    peak, peak_src = measured_peak()
    achieved = b_alg / (emit_ms * 1e-3) / GB
    path_achieved = b_alg / ((emit_ms + count_ms + scan_ms) * 1e-3) / GB
    roofline = {"bound": "hbm", "kernel": "WalkText<CSV,uint32,EMIT>", "achieved": achieved,
                "peak": peak, "unit": "GB/s", "frac": achieved / peak,
                "traffic": ncu_traffic("WalkText<2, unsigned int, 1>"),
                "peak_source": peak_src, "algorithmic_bytes_per_launch": b_alg,
                "launch_ms": emit_ms,
                "all_kernels": {"count_ms": count_ms, "scan_ms": scan_ms, "emit_ms": emit_ms,
                                "achieved": path_achieved, "frac": path_achieved / peak}}

    # ---------------- the link the end-to-end number rides on -------------------------------------
    link = None
    if not args.no_e2e:
        nb = min(n, 512 << 20)
        d_tmp = torch.empty(nb, dtype=torch.uint8, device="cuda")
        h_out = torch.empty(nb, dtype=torch.uint8, pin_memory=True)
        s_up, s_dn = torch.cuda.Stream(), torch.cuda.Stream()

        def link_rate(up, down, reps=4):
            torch.cuda.synchronize()
            t0 = time.perf_counter()
            for _ in range(reps):
                if up:
                    with torch.cuda.stream(s_up):
                        d_tmp.copy_(host[:nb], non_blocking=True)
                if down:
                    with torch.cuda.stream(s_dn):
                        h_out.copy_(dev[:nb], non_blocking=True)
            torch.cuda.synchronize()
            return reps * nb / (time.perf_counter() - t0) / GB

        link_rate(True, True, 1)
        link = {"h2d_only_GBps": link_rate(True, False), "d2h_only_GBps": link_rate(False, True),
                "h2d_with_concurrent_d2h_GBps": link_rate(True, True),
                "how": "pinned cudaMemcpyAsync of %d MiB blocks, 4 reps" % (nb >> 20)}
        del d_tmp, h_out

    # ---------------- end to end: pinned host text -> pinned host CSR -------------------------
    e2e = None
    if not args.no_e2e:
        pe = Parser("csv", index_bytes=4, label_column=LABEL_COLUMN, device=local,
                    range_bytes=args.range_bytes, chunk_bytes=args.chunk_mb << 20, depth=args.depth)
        base_ptr = harr.ctypes.data
        if args.host_threads >= 0:
            pe.set_host_threads(args.host_threads)

        handed = [0]   # bytes of the arrays handed to the caller (all seven, host memory)

        def take():
            r = pe.next_raw()
            handed[0] += 8 * (r.rows + 1) + 4 * r.rows + 8 * r.nnz
            return r.rows

        def e2e_step():
            got_rows = 0
            for _ in range(copies):
                for (a, b) in chunks:
                    while pe.pending() >= args.depth:
                        got_rows += take()
                    pe.submit_in_place(base_ptr + a, b - a)
            while pe.pending():
                got_rows += take()
            return got_rows

        e2e_steps = max(1, min(args.steps, 3))
        assert e2e_step() == rows * copies
        e2e_step()
        barrier()
        up0, down0 = pe.copied_bytes()   # the library's own count of what crossed the link
        handed[0] = 0
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e2e_step()
        torch.cuda.synchronize()
        e2e_s = max_over_ranks(time.perf_counter() - t0) / e2e_steps
        up1, down1 = pe.copied_bytes()
        barrier()
        e2e = {"value": world * text_bytes / e2e_s / GB, "unit": "GB/s",
               "h2d_bytes_per_step": (up1 - up0) // e2e_steps,
               "d2h_bytes_per_step": (down1 - down0) // e2e_steps,
               "result_bytes_per_step": handed[0] // e2e_steps,
               "index_rebuilt_on_host": (down1 - down0) < handed[0],
               "host_threads": args.host_threads if args.host_threads >= 0 else "library default (<= 4)",
               "ms_per_step": e2e_s * 1e3, "chunk_bytes": args.chunk_mb << 20, "depth": args.depth,
               "api": "b200p_submit / b200p_next (pinned host text in, pinned host CSR out)",
               "link": link,
               "frac_of_h2d_link_under_duplex": (text_bytes / e2e_s / GB) / link["h2d_with_concurrent_d2h_GBps"]}
        pe.close()

    # ---------------- CPU baseline (rank 0, N == 1) -------------------------------------------
    cpu = None
    if rank == 0 and world == 1 and not args.no_cpu_baseline:
        sample_bytes = min(n, 500 * 1000 * 1000)
        sample_bytes = chunks_end_at(harr, sample_bytes)
        path = "/dev/shm/b200_bench_sample.csv" if os.path.isdir("/dev/shm") else "/tmp/b200_bench_sample.csv"
        harr[:sample_bytes].tofile(path)
        nthread = os.cpu_count() or 1
        res = time_reference(path, nthread, args.cpu_seconds)
        one = time_reference(path, 1, min(4.0, args.cpu_seconds)) if res is not None and nthread > 1 else None
        os.remove(path)
        if res is not None:
            cpu = {"value": res["bytes"] / res["per_pass"] / GB, "unit": "GB/s", "cores": nthread,
                   "kind": "reference",
                   "sample": "first %d MB of the workload, %.1f s of passes, OMP threads=%d"
                             % (sample_bytes // 1000000, args.cpu_seconds, nthread),
                   "rows_per_s": res["rows"] / res["per_pass"]}
            if one is not None:   # OMP_NUM_THREADS in {1, all cores}
                cpu["single_thread_GBps"] = one["bytes"] / one["per_pass"] / GB

    if rank == 0:
        line = {
            "metric": "parsed input GB/s", "value": value, "unit": "GB/s", "n_gpus": world,
            "steps": args.steps, "warmup": max(args.warmup, 3), "ms_per_step": dev_ms,
            "higher_is_better": True, "scaling": "weak", "vs_baseline": None, "dtype": "f32",
            "data": "synthetic (seeded generator; one unique %d MB block tiled %dx per GPU)"
                    % (n // 1000000, copies),
            "config": {"workload": "dense CSV parse, 10 GB synthetic, 128 float columns + label_column=0, "
                                   "Parser<uint32_t>, 1xB200 per rank (BASELINE.json configs[1])",
                       "text_bytes_per_gpu": text_bytes, "rows_per_gpu": rows * copies,
                       "nnz_per_gpu": nnz * copies, "window_bytes": n,
                       "l2": "inputs (10 GB) far larger than the 126 MB L2; no flush needed",
                       "parallelism": "shard-per-GPU, no collective",
                       "streams": len(lanes)},
            "rows_per_s": world * rows * copies / (dev_ms * 1e-3),
            "gpu_launches": int(launches),
            "clocks": clocks,
            "roofline": roofline,
            "setup_s": setup_s,
        }
        if e2e is not None:
            line["e2e"] = e2e
        if cpu is not None:
            line["cpu_baseline"] = cpu
        print(json.dumps(line))
    for (q, _) in lanes:
        q.close()
    if world > 1:
        dist.destroy_process_group()


def chunks_end_at(arr, limit):
    """largest prefix <= limit that ends right after a newline"""
    back = arr[:limit][::-1]
    return limit - int(np.argmax(back == 10))


if __name__ == "__main__":
    main()
