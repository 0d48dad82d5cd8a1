"""Per-stage counters of the extractor from an ncu full-set summary (tools/ncu_summary.py CSV): executed warp
instructions, DRAM bytes and the busiest pipe per frame.  bench.py turns them into issue / pipe fractions with the
LIVE stage times (instruction counts do not depend on the clock or on what else runs; times under ncu do).
    python tools/ncu_counters.py profiles/r2_ncu_full_extract.csv 256 profiles/r2_kernel_counters.json
"""
import csv
import json
import sys

STAGE = {"resize": "pyramid", "fast_cells": "fast", "select": "select", "blur7": "blur", "describe": "describe"}


def main(src, frames_per_launch, out):
    rows = list(csv.reader(open(src)))
    h = rows[0]
    col = {n: h.index(n) for n in h}
    acc = {}
    nl = {}
    for r in rows[2:]:
        st = next((v for k, v in STAGE.items() if k in r[0]), None)
        if not st:
            continue
        a = acc.setdefault(st, {"kernel": r[0].split("(")[0].replace("void ", ""), "inst": 0.0, "dram": 0.0, "us": 0.0,
                                "alu": 0.0, "fma": 0.0, "lsu": 0.0, "xu": 0.0, "issue": 0.0, "warps": 0.0})
        t = float(r[col["gpu__time_duration.sum"]])
        a["inst"] += float(r[col["smsp__inst_executed.sum"]])
        a["dram"] += (float(r[col["dram__bytes_read.sum"]]) + float(r[col["dram__bytes_write.sum"]])) * 1e6
        a["us"] += t
        for k, m in (("alu", "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active"),
                     ("fma", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active"),
                     ("lsu", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active"),
                     ("xu", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active"),
                     ("issue", "smsp__issue_active.avg.pct_of_peak_sustained_active"),
                     ("warps", "sm__warps_active.avg.pct_of_peak_sustained_active")):
            a[k] += float(r[col[m]]) * t                       # time-weighted
        nl[st] = nl.get(st, 0) + 1
    passes = max(nl.get("fast", 1), 1)                         # chunks captured
    res = {}
    for st, a in acc.items():
        f = frames_per_launch * passes
        pipes = {k: a[k] / a["us"] for k in ("alu", "fma", "lsu", "xu")}
        top = max(pipes, key=pipes.get)
        res[st] = {"kernel": a["kernel"], "warp_inst_per_frame": a["inst"] / f, "dram_bytes_per_frame": a["dram"] / f,
                   "ncu_us_per_frame": a["us"] / f, "ncu_issue_active_pct": a["issue"] / a["us"],
                   "ncu_warps_active_pct": a["warps"] / a["us"], "busiest_pipe": top, "ncu_busiest_pipe_pct": pipes[top]}
    json.dump({"source": src, "frames_per_launch": frames_per_launch, "chunks": passes, "stages": res}, open(out, "w"), indent=1)
    for st, v in res.items():
        print(st, {k: (round(x, 2) if isinstance(x, float) else x) for k, x in v.items()})


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]), sys.argv[3])
