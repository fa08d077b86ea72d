#!/usr/bin/env python
"""Turn an .ncu-rep (ncu --set full) into the short per-launch summary kept under profiles/.

    python tools/summarize_ncu.py gpurun_out/prof.ncu-rep > profiles/rNN_<what>_ncu_full.txt
"""
import csv
import io
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "dram__cycles_active.avg.pct_of_peak_sustained_elapsed",
    "lts__t_bytes.sum", "lts__t_sector_hit_rate.pct", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__warps_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_tensor.sum", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_fma.sum", "sm__inst_executed_pipe_alu.sum", "sm__inst_executed_pipe_lsu.sum",
    "smsp__inst_executed.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "launch__registers_per_thread", "launch__block_size",
    "launch__grid_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem",
    "launch__occupancy_limit_registers", "launch__waves_per_multiprocessor", "sm__cycles_elapsed.avg",
    "sm__cycles_active.avg", "smsp__cycles_active.avg",
]


def main():
    rep = sys.argv[1]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units = rows[0], rows[1]
    col = {h: i for i, h in enumerate(hdr)}
    print(f"# {rep}: ncu --set full --clock-control none, one block per captured launch")
    for n, r in enumerate(rows[2:]):
        print(f"\n## launch {n}: {r[col['Kernel Name']]}")
        for k in KEYS:
            if k in col and r[col[k]] != "":
                print(f"{k:70s} {r[col[k]]:>18s} {units[col[k]]}")
        stalls = [(float(r[i] or 0), h) for h, i in col.items()
                  if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
        if not stalls:
            stalls = [(float(r[i] or 0), h) for h, i in col.items()
                      if h.startswith("smsp__average_warp_latency_issue_stalled_") and h.endswith(".ratio")]
        for v, h in sorted(stalls, reverse=True)[:6]:
            print(f"{h:70s} {v:18.3f}")


if __name__ == "__main__":
    main()
