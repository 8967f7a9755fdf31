"""Summarise an .ncu-rep (read with `ncu -i … --page raw --csv`) into the handful of counters the roofline needs.

    python tools/ncu_summary.py gpurun_out/prof.ncu-rep [more metric substrings…]
"""
import csv
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "gpu__dram_throughput.avg.pct",
        "dram__throughput.avg.pct", "sm__throughput.avg.pct", "sm__warps_active.avg.pct", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "launch__occupancy_limit", "sm__inst_executed.sum ",
        "smsp__inst_executed.avg.per_cycle_active", "sm__inst_executed.avg.per_cycle_elapsed",
        "sm__pipe_tensor_cycles_active", "sm__pipe_tensor_subunit", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared",
        "smsp__inst_executed_op_local", "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "lts__t_bytes.sum ", "l1tex__t_bytes.sum ", "smsp__average_warp", "smsp__warp_issue_stalled",
        "smsp__average_warps_issue_stalled", "sm__cycles_elapsed.max", "l1tex__throughput.avg.pct",
        "lts__throughput.avg.pct", "smsp__pcsamp_warps_issue_stalled"]


def main() -> None:
    rep = sys.argv[1]
    extra = sys.argv[2:]
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        print("=" * 100)
        print(r[hdr.index("Kernel Name")][:120], "grid", r[hdr.index("Grid Size")], "block", r[hdr.index("Block Size")])
        for i, h in enumerate(hdr):
            name = h + " "
            if any(k in name for k in KEYS + extra):
                v = r[i]
                if v in ("", "0", "n/a"):
                    continue
                print(f"  {h:100s} {v} {units[i]}")


if __name__ == "__main__":
    main()
