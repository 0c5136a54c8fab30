#!/usr/bin/env python
"""Turn an ncu report brought back in gpurun_out/ into the small tracked summaries under
profiles/:   python profiles/summarize_ncu.py gpurun_out/<name>.ncu-rep profiles/<out-stem> [text_bytes]
writes <out-stem>.metrics.json (per-kernel headline metrics incl. DRAM bytes per launch; text_bytes,
if given, records the size of the chunk the launches parsed so that bench.py can scale the traffic
to its own window) and <out-stem>.opcodes.txt (dynamic SASS opcode mix, ALU-pipe vs FMA-pipe share)."""
import collections
import csv
import io
import json
import re
import subprocess
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size",
        "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__thread_inst_executed_per_inst_executed.ratio",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio"]
ALU = {"LOP3", "IADD3", "SHF", "PRMT", "ISETP", "SEL", "LEA", "IABS", "IMNMX", "VIMNMX", "FSEL",
       "FSETP", "PLOP3", "LOP", "IADD", "MOV", "SGXT", "BMSK", "FMNMX", "VIADD", "VIADDMNMX",
       "P2R", "R2P"}
FMA = {"IMAD", "FFMA", "FMUL", "FADD", "HFMA2", "IDP", "IDP4A"}


def ncu(args):
    return subprocess.run(["ncu"] + args, stdout=subprocess.PIPE, stderr=subprocess.DEVNULL,
                          text=True).stdout


def unit_scale(unit):
    return {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6,
            "ns": 1e-9, "s": 1.0}.get(unit, None)


def main():
    rep, stem = sys.argv[1], sys.argv[2]
    text_bytes = int(sys.argv[3]) if len(sys.argv) > 3 else None
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "raw", "--csv"]))))
    hdr, units, data = rows[0], rows[1], rows[2:]
    out = []
    for r in data:
        k = {"kernel": r[hdr.index("Kernel Name")], "capture": stem.split("/")[-1]}
        if text_bytes is not None:
            k["text_bytes"] = text_bytes
        for m in WANT:
            if m in hdr:
                i = hdr.index(m)
                try:
                    v = float(r[i].replace(",", ""))
                except ValueError:
                    continue
                sc = unit_scale(units[i])
                if m.startswith("dram__bytes") or m.startswith("gpu__time"):
                    v = v * (sc or 1.0)
                k[m] = v
        if "dram__bytes_read.sum" in k:
            k["dram_bytes_per_launch"] = k["dram__bytes_read.sum"] + k.get("dram__bytes_write.sum", 0.0)
        out.append(k)
    with open(stem + ".metrics.json", "w") as f:
        json.dump(out, f, indent=1)
    rows = list(csv.reader(io.StringIO(ncu(["-i", rep, "--page", "source", "--csv",
                                            "--print-source", "sass"]))))
    lines, i, seen = [], 0, set()
    while i < len(rows):
        r = rows[i]
        if r and r[0] == "Kernel Name" and r[1] not in seen:
            seen.add(r[1])
            h = rows[i + 1]
            si, ie = h.index("Source"), h.index("Instructions Executed")
            hist, tot, alu, fma = collections.Counter(), 0, 0, 0
            i += 2
            while i < len(rows) and rows[i] and rows[i][0] != "Kernel Name":
                try:
                    n = int(rows[i][ie])
                except (ValueError, IndexError):
                    i += 1
                    continue
                m = re.match(r"\s*(?:@!?U?P\d+\s+)?([A-Z0-9_.]+)", rows[i][si])
                op = m.group(1).split(".")[0] if m else "?"
                hist[op] += n
                tot += n
                alu += n if op in ALU else 0
                fma += n if op in FMA else 0
                i += 1
            lines.append("%s\n  warp instructions %d  ALU-pipe %.1f%%  FMA-pipe %.1f%%\n  %s\n" % (
                r[1], tot, 100.0 * alu / max(tot, 1), 100.0 * fma / max(tot, 1),
                "  ".join("%s %.1f%%" % (k, 100.0 * n / max(tot, 1)) for k, n in hist.most_common(20))))
            continue
        i += 1
    with open(stem + ".opcodes.txt", "w") as f:
        f.write("\n".join(lines))
    print(json.dumps(out, indent=1)[:1500])


if __name__ == "__main__":
    main()
