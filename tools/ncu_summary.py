"""Markdown summary of an `ncu --set full` report:  python tools/ncu_summary.py report.ncu-rep > profiles/x.md
One block per captured launch: duration, DRAM traffic, pipe utilisation, occupancy, top stall reasons."""
import csv
import io
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "duration"),
    ("dram__bytes_read.sum", "DRAM read"),
    ("dram__bytes_write.sum", "DRAM write"),
    ("dram__throughput.avg.pct_of_peak_sustained_elapsed", "DRAM throughput % of peak"),
    ("lts__t_sector_hit_rate.pct", "L2 hit rate %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active", "FP64 pipe % (inst, active)"),
    ("sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "FP64 pipe cycles active %"),
    ("sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "LSU pipe %"),
    ("sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "ALU pipe %"),
    ("sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "FMA pipe %"),
    ("l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed", "L1 LSU wavefronts %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("sm__warps_active.avg.pct_of_peak_sustained_active", "achieved occupancy %"),
    ("launch__occupancy_limit_registers", "occupancy limit (registers), blocks/SM"),
    ("launch__occupancy_limit_shared_mem", "occupancy limit (shared mem), blocks/SM"),
    ("launch__registers_per_thread", "registers/thread"),
    ("launch__shared_mem_per_block_dynamic", "dynamic smem/block"),
    ("smsp__inst_executed.sum", "warp instructions"),
    ("smsp__sass_thread_inst_executed_op_dfma_pred_on.sum", "DFMA thread inst"),
    ("smsp__sass_thread_inst_executed_op_dmul_pred_on.sum", "DMUL thread inst"),
    ("smsp__sass_thread_inst_executed_op_dadd_pred_on.sum", "DADD thread inst"),
    ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "shared bank conflicts"),
]


def main(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units, data = rows[0], rows[1], rows[2:]
    col = {k: i for i, k in enumerate(head)}
    print(f"# {path.split('/')[-1]} — `ncu --set full --clock-control none` (per-launch, cold-ish cache, serialised)\n")
    for r in data:
        name = r[col["Kernel Name"]]
        print(f"## `{name}`  grid {r[col['Grid Size']]} block {r[col['Block Size']]}\n")
        print("| metric | value | unit |\n|---|---:|---|")
        for k, label in KEYS:
            if k in col and r[col[k]] != "":
                print(f"| {label} (`{k}`) | {r[col[k]]} | {units[col[k]]} |")
        stalls = []
        for k, i in col.items():
            if k.startswith("smsp__average_warps_issue_stalled_") and k.endswith("_per_issue_active.ratio") and r[i]:
                try:
                    stalls.append((float(r[i].replace(",", "")), k[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")]))
                except ValueError:
                    pass
        if stalls:
            stalls.sort(reverse=True)
            print("\nTop stall reasons (warps stalled per issue-active cycle): " + ", ".join(f"{n} {v:.2f}" for v, n in stalls[:5]))
        print()


if __name__ == "__main__":
    main(sys.argv[1])
