#!/usr/bin/env python
"""Summarise ncu outputs into the text files kept under profiles/.

  python profiles/summarize_ncu.py launches <launches.csv>
  python profiles/summarize_ncu.py raw <report.ncu-rep>      (needs ncu on PATH)
  python profiles/summarize_ncu.py json-score <report.ncu-rep> <events of the profiled launch>   -> profiles/score_ncu.json
  python profiles/summarize_ncu.py json <report.ncu-rep> <word visits of the profiled launch> [kernel substring]
      -> profiles/estep_ncu.json: what bench.py's roofline record reads (DRAM bytes and FP64
         lane-operations per word visit of the main E-step launch), stamped with a hash of the
         kernel sources so that bench.py refuses it once those change
"""
import hashlib
import json
import os
import csv
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "launch__grid_size", "launch__registers_per_thread",
    "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
    "sm__cycles_active.avg", "sm__cycles_active.min", "sm__cycles_active.max", "sm__cycles_elapsed.avg",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__warps_active.avg.per_cycle_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__thread_inst_executed_per_inst_executed.ratio", "l1tex__t_sector_hit_rate.pct",
    "lts__t_sector_hit_rate.pct", "sm__warps_active.avg.pct_of_peak_sustained_active",
]


def launches(path):
    rows = list(csv.reader(open(path)))
    hdr, agg = None, {}
    for r in rows:
        if len(r) > 5 and r[0] == "ID":
            hdr = r
            continue
        if hdr and len(r) == len(hdr):
            d = dict(zip(hdr, r))
            if d.get("Metric Name") == "gpu__time_duration.sum":
                v = float(d["Metric Value"].replace(",", ""))
                v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(d["Metric Unit"], 1.0)
                agg.setdefault(d["Kernel Name"].split("(")[0][:64], []).append(v)
    tot = sum(sum(v) for v in agg.values())
    print("%-66s %5s %12s %7s" % ("kernel", "n", "total ms", "share"))
    for k, v in sorted(agg.items(), key=lambda kv: -sum(kv[1])):
        print("%-66s %5d %12.3f %6.2f%%" % (k, len(v), sum(v), 100 * sum(v) / tot))


def raw(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for vals in rows[2:]:
        d = dict(zip(hdr, vals))
        print("== %s  grid %s  block %s" % (d.get("Kernel Name", "?")[:90], d.get("Grid Size"), d.get("Block Size")))
        for h, u, v in zip(hdr, units, vals):
            if h in KEYS:
                print("   %-72s %-10s %s" % (h, u, v))
        stalls = []
        for h, v in zip(hdr, vals):
            if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio"):
                try:
                    stalls.append((float(v), h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        print("   stalls per issue: " + ", ".join("%s %.2f" % (n, x) for x, n in sorted(stalls, reverse=True)[:6]))


def source_hash(names=("estep.cuh", "devmath.cuh", "math_tables.h", "estep_inst.inc")):  # = bench.py kernel_source_hash()
    root = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
    h = hashlib.sha256()
    for name in names:
        with open(os.path.join(root, "oni_ml_b200", "csrc", name), "rb") as f:
            h.update(f.read())
    return h.hexdigest()[:16]


def to_json_score(path, events, kernel="score_filter_paired_kernel"):
    """profiles/score_ncu.json: DRAM bytes per event of the scoring launch."""
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], dict(zip(rows[0], rows[1]))
    pick = [dict(zip(hdr, r)) for r in rows[2:] if kernel in dict(zip(hdr, r)).get("Kernel Name", "")]
    if len(pick) != 1:
        raise SystemExit("%d launches match %r in %s; capture exactly one" % (len(pick), kernel, path))
    d = pick[0]

    def val(name):
        v = float(d[name].replace(",", ""))
        return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1.0,
                    "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units[name], 1.0)
    dram = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    rec = {"kernel": d["Kernel Name"], "report": os.path.basename(path),
           "source_hash": source_hash(("scoring.cu", "radix_sort.cuh")), "launch_events": events,
           "duration_ms": val("gpu__time_duration.sum"), "dram_bytes": dram, "dram_bytes_per_event": dram / events,
           "dram_pct_of_peak": float(d["gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed"]),
           "l2_hit_pct": float(d["lts__t_sector_hit_rate.pct"]), "registers": int(d["launch__registers_per_thread"])}
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "score_ncu.json")
    with open(dst, "w") as f:
        json.dump(rec, f, indent=1)
        f.write("\n")
    print(json.dumps(rec, indent=1))


def to_json(path, visits, kernel="estep_kernel<4, 5"):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr = rows[0]
    pick = [dict(zip(hdr, r)) for r in rows[2:] if kernel in dict(zip(hdr, r)).get("Kernel Name", "")]
    if len(pick) != 1:
        raise SystemExit("%d launches match %r in %s; capture exactly one (-k ... -c 1)" % (len(pick), kernel, path))
    d = pick[0]
    units = dict(zip(hdr, rows[1]))

    def val(name):
        v = float(d[name].replace(",", ""))
        return v * {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "byte": 1.0,
                    "ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(units[name], 1.0)
    dram = val("dram__bytes_read.sum") + val("dram__bytes_write.sum")
    # FP64 lane-operations: thread-level executions of every D* arithmetic instruction (source page)
    src = subprocess.run(["ncu", "-i", path, "--page", "source", "--csv"], capture_output=True, text=True).stdout
    srows = list(csv.reader(src.splitlines()))
    h2 = next(r for r in srows if "Source" in r and "Thread Instructions Executed" in r)
    i_src, i_thr, i_ins = h2.index("Source"), h2.index("Thread Instructions Executed"), h2.index("Instructions Executed")
    lane_ops = warp_inst = fp64_inst = 0
    for r in srows[srows.index(h2) + 1:]:
        if len(r) <= i_thr or not r[i_thr].isdigit():
            continue
        op = r[i_src].split()
        op = op[1] if op and op[0].startswith("@") else (op[0] if op else "")
        warp_inst += int(r[i_ins])
        if op.split(".")[0] in ("DFMA", "DMUL", "DADD"):
            lane_ops += int(r[i_thr])
            fp64_inst += int(r[i_ins])
    rec = {
        "kernel": d["Kernel Name"], "report": os.path.basename(path), "source_hash": source_hash(),
        "launch_word_visits": visits, "duration_ms": val("gpu__time_duration.sum"),
        "dram_bytes": dram, "dram_bytes_per_visit": dram / visits,
        "fp64_lane_ops": lane_ops, "fp64_lane_ops_per_visit": lane_ops / visits,
        "fp64_warp_instructions": fp64_inst, "warp_instructions": warp_inst,
        "fp64_pipe_pct": float(d["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]),
        "issue_active_pct": float(d["smsp__issue_active.avg.pct_of_peak_sustained_active"]),
        "l2_hit_pct": float(d["lts__t_sector_hit_rate.pct"]),
        "registers": int(d["launch__registers_per_thread"]),
    }
    dst = os.path.join(os.path.dirname(os.path.abspath(__file__)), "estep_ncu.json")
    with open(dst, "w") as f:
        json.dump(rec, f, indent=1)
        f.write("\n")
    print(json.dumps(rec, indent=1))


if __name__ == "__main__":
    if sys.argv[1] == "json":
        to_json(sys.argv[2], float(sys.argv[3]), *sys.argv[4:5])
    elif sys.argv[1] == "json-score":
        to_json_score(sys.argv[2], float(sys.argv[3]), *sys.argv[4:5])
    else:
        {"launches": launches, "raw": raw}[sys.argv[1]](sys.argv[2])
