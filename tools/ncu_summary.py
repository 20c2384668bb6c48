#!/usr/bin/env python
"""Condense an `ncu --set full` report into the per-kernel table kept under profiles/.
Usage: python tools/ncu_summary.py report.ncu-rep out.csv   (needs the `ncu` CLI; reads the report, runs nothing)"""
import csv
import subprocess
import sys

KEEP = [
    ("Kernel Name", "kernel"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"), ("launch__occupancy_limit_registers", "occ_limit_regs_blocks"),
    ("launch__occupancy_limit_shared_mem", "occ_limit_smem_blocks"), ("gpu__time_duration.sum", "duration_us"),
    ("dram__bytes_read.sum", "dram_read"), ("dram__bytes_write.sum", "dram_write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "dram_pct_of_peak"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm_pct_of_peak"),
    ("smsp__inst_executed.sum", "warp_inst"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue_active_pct"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active", "fma_pipe_active_pct"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "alu_pipe_pct"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "xu_pipe_pct"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved_occupancy_pct"),
    ("smsp__warps_eligible.avg.per_cycle_active", "eligible_warps_per_cycle"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem_bank_conflicts"),
]


def main():
    rep, out = sys.argv[1], sys.argv[2]
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True, check=True).stdout
    rows = list(csv.reader(raw.splitlines()))
    hdr, units, data = rows[0], rows[1], rows[2:]
    stall = [(i, h.replace("smsp__average_warps_issue_stalled_", "").replace("_per_issue_active.ratio", ""))
             for i, h in enumerate(hdr) if "issue_stalled" in h and h.endswith("per_issue_active.ratio")]
    with open(out, "w", newline="") as f:
        w = csv.writer(f)
        w.writerow([short + (f" [{units[hdr.index(k)]}]" if units[hdr.index(k)] else "") for k, short in KEEP
                    if k in hdr] + ["top_stalls(cycles per issue)"])
        for d in data:
            top = sorted(((float(d[i]), n) for i, n in stall if d[i] not in ("", "n/a")), reverse=True)[:5]
            w.writerow([d[hdr.index(k)] for k, _ in KEEP if k in hdr] + [" ".join(f"{n}={v:.2f}" for v, n in top)])
    print(out)


if __name__ == "__main__":
    main()
