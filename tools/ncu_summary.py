#!/usr/bin/env python
"""Summarise an .ncu-rep (read here, no GPU needed) into the few numbers DESIGN.md and
bench.py's roofline cite.  Usage: tools/ncu_summary.py <rep> [<rep> ...] > profiles/<name>.md"""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_elapsed", "ALU pipe active % (elapsed)"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe inst % (active)"),
    ("sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed", "FMA pipe active %"),
    ("sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_elapsed", "XU pipe %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_elapsed", "LSU pipe %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots active %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__registers_per_thread", "registers / thread"),
    ("launch__grid_size", "grid"),
    ("launch__block_size", "block"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit: shared mem (blocks)"),
    ("launch__occupancy_limit_registers", "occupancy limit: registers (blocks)"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts"),
    ("l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "smem wavefronts"),
]
STALL = "smsp__average_warps_issue_stalled_"


def main():
    for rep in sys.argv[1:]:
        raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
        rows = list(csv.reader(io.StringIO(raw)))
        hdr, units, vals = rows[0], rows[1], rows[2:]
        print("## %s\n" % rep.split("/")[-1])
        for v in vals:
            name = v[hdr.index("Kernel Name")]
            print("### `%s`\n" % name[:110])
            print("| metric | value |\n|---|---|")
            col = {h: (x, u) for h, u, x in zip(hdr, units, v)}
            for k, label in KEYS:
                if k in col:
                    print("| %s (`%s`) | %s %s |" % (label, k, col[k][0], col[k][1]))
            if "dram__bytes_read.sum" in col:
                def to_b(x, u):
                    return float(x) * {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
                tr = to_b(*col["dram__bytes_read.sum"]) + to_b(*col["dram__bytes_write.sum"])
                print("| **traffic = DRAM read + write** | %.0f bytes |" % tr)
            st = sorted(((float(col[h][0]), h[len(STALL):-len("_per_issue_active.ratio")]) for h in col
                         if h.startswith(STALL) and h.endswith("_per_issue_active.ratio")), reverse=True)
            print("| top stall reasons (warps per issue) | %s |" % ", ".join("%s %.2f" % (n, x) for x, n in st[:6]))
            print()


if __name__ == "__main__":
    main()
