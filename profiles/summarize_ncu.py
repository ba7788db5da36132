"""Extracts the judged metrics from an .ncu-rep (read on the CPU box with `ncu -i`).
usage: python profiles/summarize_ncu.py gpurun_out/prof.ncu-rep > profiles/rNN_xxx.txt"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__occupancy_limit_registers", "smsp__inst_executed.sum",
        "smsp__thread_inst_executed_per_inst_executed.ratio", "smsp__sass_average_branch_targets_threads_uniform.pct",
        "smsp__warps_eligible.avg.per_cycle_active", "smsp__cycles_active.avg"]


def main():
    out = subprocess.run(["ncu", "-i", sys.argv[1], "--page", "raw", "--csv"], stdout=subprocess.PIPE, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    H, U, data = rows[0], rows[1], rows[2:]
    ki = H.index("Kernel Name")
    print("metric,unit," + ",".join(r[ki].split("(")[0].split("::")[-1] for r in data))
    for k in KEYS:
        if k in H:
            i = H.index(k)
            print(f"{k},{U[i]}," + ",".join(r[i] for r in data))
    for i, h in enumerate(H):
        if "warp_issue_stalled" in h and h.endswith("per_warp_active.pct"):
            vals = [float(r[i].replace(",", "")) for r in data]
            if max(vals) >= 3:
                print(f"{h},{U[i]}," + ",".join(f"{v:.1f}" for v in vals))


if __name__ == "__main__":
    main()
