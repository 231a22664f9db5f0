#!/usr/bin/env python
"""profiles/rNN/ncu_dominant_kernel.json from the raw page of one `ncu --set full` capture:

  ncu -i X.ncu-rep --page raw --csv > X.csv
  python tools/ncu_summary.py X.csv SHOTS_PER_LAUNCH "description" > ncu_dominant_kernel.json

bench.py reads `dram_bytes_per_shot` from the result for `roofline.traffic`."""
import csv
import json
import re
import sys


def main():
    rows = list(csv.reader(open(sys.argv[1])))
    shots = int(sys.argv[2])
    hdr, units, vals = rows[0], rows[1], rows[2]
    d = dict(zip(hdr, vals))
    u = dict(zip(hdr, units))

    def f(k):
        return float(d[k].replace(",", ""))

    def nbytes(k):
        return f(k) * {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[u[k]]

    stalls = {h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""): round(f(h), 2)
              for h in hdr if re.search(r"smsp__average_warps?_issue_stalled_.*_per_issue_active", h)
              and d[h] not in ("", "n/a")}
    stalls = dict(sorted(stalls.items(), key=lambda kv: -kv[1])[:6])
    t_ms = f("gpu__time_duration.sum") * {"ms": 1, "us": 1e-3, "s": 1e3, "ns": 1e-6}[u["gpu__time_duration.sum"]]
    out = {
        "kernel": d.get("Kernel Name", "") + " -- " + (sys.argv[3] if len(sys.argv) > 3 else ""),
        "capture": "ncu --set full --clock-control none --import-source on -k regex:k_sample_l -s 3 -c 1 python "
                   "bench.py --steps 2 --warmup 3 --no-cpu-baseline --no-enum (profiles/profile_cmd.sh)",
        "variant_key": sys.argv[4] if len(sys.argv) > 4 else "",
        "shots_per_launch": shots,
        "gpu__time_duration_ms": t_ms,
        "dram__bytes_write.sum": nbytes("dram__bytes_write.sum"),
        "dram__bytes_read.sum": nbytes("dram__bytes_read.sum"),
        "dram_bytes_per_shot": (nbytes("dram__bytes_write.sum") + nbytes("dram__bytes_read.sum")) / shots,
        "warp_instructions_per_1024_shots": round(f("smsp__inst_executed.sum") / shots * 1024),
        "stalls_per_issue": stalls,
    }
    for k in ("smsp__inst_executed.sum", "smsp__issue_active.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
              "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
              "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
              "smsp__inst_executed_op_shared_atom.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
              "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread",
              "launch__block_size", "launch__grid_size", "sm__warps_active.avg.pct_of_peak_sustained_active"):
        if k in d and d[k] not in ("", "n/a"):
            out[k] = f(k)
    json.dump(out, sys.stdout, indent=1)
    print()


if __name__ == "__main__":
    main()
