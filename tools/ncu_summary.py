"""Key metrics of every kernel in one or more .ncu-rep files, as markdown.

    python tools/ncu_summary.py gpurun_out/a.ncu-rep [b.ncu-rep ...] > profiles/rNN_summary.md
"""
import csv
import io
import subprocess
import sys

WANT = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct",
    "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smsp__average_warp_latency_issue_stalled_no_instruction.ratio",
    "sm__inst_executed_pipe_fp64.sum", "smsp__warp_issue_stalled_no_instruction_per_warp_active.pct",
]
for path in sys.argv[1:]:
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(hdr, r))
        u = dict(zip(hdr, units))
        print(f"## {d['Kernel Name'][:90]}  ({path.split('/')[-1]})")
        for k in WANT:
            if k in d:
                print(f"- {k} = {d[k]} {u[k]}")
        stalls = sorted(((float(v.replace(',', '')), k) for k, v in d.items()
                         if k.startswith("smsp__average_warps_issue_stalled") and k.endswith("_per_issue_active.ratio") and v),
                        reverse=True)[:6]
        if stalls:
            print("- top stalls (warps per issue): " + ", ".join(f"{k.split('stalled_')[1].split('_per_')[0]} {v:.2f}" for v, k in stalls))
        print()
