#!/usr/bin/env python
"""Key metrics per launch of an .ncu-rep (`ncu --set full`) as a small CSV for profiles/ (no GPU needed).

  python tools/ncu_summary.py REPORT.ncu-rep > profiles/rNN_<what>.csv
"""
import csv
import subprocess
import sys

WANT = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "lts__t_sector_hit_rate.pct"]
STALL = "smsp__average_warps_issue_stalled_"


def main():
    txt = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], check=True, stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(txt.splitlines()))
    hdr, units = rows[0], rows[1]
    cols = [c for c in WANT if c in hdr]
    stalls = [h for h in hdr if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")]
    w = csv.writer(sys.stdout)
    w.writerow([f"{c} [{units[hdr.index(c)]}]" if units[hdr.index(c)] else c for c in cols] + ["top stalls (warps per issue)"])
    for r in rows[2:]:
        st = sorted(((float(r[hdr.index(h)].replace(",", "") or 0), h[len(STALL):-len("_per_issue_active.ratio")]) for h in stalls), reverse=True)
        w.writerow([r[hdr.index(c)] for c in cols] + ["; ".join(f"{n} {v:.2f}" for v, n in st[:6])])


if __name__ == "__main__":
    main()
