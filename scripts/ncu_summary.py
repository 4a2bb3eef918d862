"""Condense an .ncu-rep (read with `ncu -i ... --page raw --csv`) into the few columns the roofline uses.

    python scripts/ncu_summary.py gpurun_out/prof.ncu-rep profiles/r01_xxx.csv
"""
import csv
import io
import subprocess
import sys

KEEP = ["Kernel Name", "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
        "l1tex__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "smsp__pcsamp_warps_issue_stalled_long_scoreboard",
        "smsp__pcsamp_warps_issue_stalled_barrier", "smsp__pcsamp_warps_issue_stalled_mio_throttle",
        "smsp__pcsamp_warps_issue_stalled_lg_throttle", "smsp__pcsamp_warps_issue_stalled_math_pipe_throttle",
        "smsp__pcsamp_warps_issue_stalled_short_scoreboard", "smsp__pcsamp_warps_issue_stalled_wait",
        "smsp__pcsamp_warps_issue_stalled_selected", "smsp__pcsamp_sample_buffer_full"]


def main(rep, out):
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(io.StringIO(raw)))
    hdr, units, data = rows[0], rows[1], rows[2:]
    idx = [hdr.index(k) for k in KEEP if k in hdr]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([hdr[i] for i in idx])
        w.writerow([units[i] for i in idx])
        for r in data:
            w.writerow([r[i] for i in idx])
    for r in data:
        d = {hdr[i]: r[i] for i in idx}
        tr = float(d["dram__bytes_read.sum"]) + float(d["dram__bytes_write.sum"])
        print(f'{d["Kernel Name"][:60]:60s} {float(d["gpu__time_duration.sum"]):9.1f} {units[hdr.index("gpu__time_duration.sum")]} '
              f'dram {tr:9.2f} {units[hdr.index("dram__bytes_read.sum")]}  regs {d["launch__registers_per_thread"]} '
              f'warps_active {float(d["sm__warps_active.avg.pct_of_peak_sustained_active"]):5.1f}% '
              f'fp64 {float(d["sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active"]):5.1f}% '
              f'L2hit {float(d["lts__t_sector_hit_rate.pct"]):5.1f}% L1hit {float(d["l1tex__t_sector_hit_rate.pct"]):5.1f}%')


if __name__ == "__main__":
    main(sys.argv[1], sys.argv[2])
