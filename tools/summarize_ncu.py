#!/usr/bin/env python
"""Summarise an Nsight Compute report (gpurun_out/*.ncu-rep) into profiles/.

    python tools/summarize_ncu.py gpurun_out/prof.ncu-rep <tag> [workload]

Writes profiles/<tag>_ncu_summary.txt (selected raw metrics per captured kernel) and merges the per-launch DRAM
traffic of every kernel into profiles/ncu_summary.json, which bench.py reads for `roofline.traffic`.
"""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "launch__grid_size", "launch__block_size",
        "launch__registers_per_thread", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_tensor_cycles_active_realtime.avg.pct_of_peak_sustained_elapsed", "smsp__pipe_tensor_subpipe_dmma_cycles_active.avg",
        "sm__cycles_elapsed.avg", "smsp__inst_executed.sum", "lts__t_sector_hit_rate.pct", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "smsp__sass_thread_inst_executed_op_dmul_pred_on.sum",
        "smsp__sass_thread_inst_executed_op_dadd_pred_on.sum",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio", "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio", "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio", "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio"]
KNAME = {"k_geom_coop": "k_geom_coop/k_geom", "k_geom": "k_geom_coop/k_geom"}


def main():
    rep, tag = sys.argv[1], sys.argv[2]
    workload = sys.argv[3] if len(sys.argv) > 3 else "big_tricylinder"
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    ik = hdr.index("Kernel Name")
    out = [f"# {os.path.basename(rep)}  ({workload}; ncu --set full --clock-control none; per-launch values, cold-cache replay)"]
    summ_path = os.path.join(ROOT, "profiles", "ncu_summary.json")
    summ = json.load(open(summ_path)) if os.path.exists(summ_path) else {}
    for r in data:
        name = r[ik].split("(")[0].replace("void ", "").split("<")[0]
        out.append(f"\n## {r[ik]}")
        vals = {}
        for i, h in enumerate(hdr):
            base = h.split(".TriageCompute.")[-1] if ".TriageCompute." in h else h
            if base in KEEP and r[i] != "":
                out.append(f"{base} [{units[i]}] = {r[i]}")
                try:
                    vals[base] = (float(r[i].replace(",", "")), units[i])
                except ValueError:                              # "no data" for a metric ncu could not collect
                    pass
        def to_bytes(key):
            v, u = vals.get(key, (0.0, "byte"))
            return v * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
        dram = to_bytes("dram__bytes_read.sum") + to_bytes("dram__bytes_write.sum")
        if dram == dram and abs(dram) != float("inf"):      # kernels ncu could not replay report NaN: keep them out of the JSON
            summ.setdefault(workload, {})[KNAME.get(name, name)] = {"dram_bytes_per_launch": dram, "source": os.path.basename(rep), "tag": tag}
    with open(os.path.join(ROOT, "profiles", f"{tag}_ncu_summary.txt"), "w") as fh:
        fh.write("\n".join(out) + "\n")
    with open(summ_path, "w") as fh:
        json.dump(summ, fh, indent=1, sort_keys=True)
    print("\n".join(out[:60]))


if __name__ == "__main__":
    main()
