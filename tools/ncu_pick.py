"""Key metrics of `ncu -i X.ncu-rep --page raw --csv` files (one kernel per file), as a small table.
usage: python tools/ncu_pick.py a_raw.csv [b_raw.csv ...] [--csv out.csv]"""
import csv
import sys

WANT = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "smsp__inst_executed.sum",
        "sm__inst_executed.avg.per_cycle_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active",
        "sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
        "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
        "l1tex__t_requests_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
        "l1tex__t_output_wavefronts_pipe_lsu_mem_global_op_ld.sum", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "l1tex__throughput.avg.pct_of_peak_sustained_active", "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sectors_srcunit_tex_op_read.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
        "launch__shared_mem_per_block_dynamic", "sm__cycles_active.avg", "sm__cycles_elapsed.max"]


def load(path):
    rows = list(csv.reader(open(path)))
    hdr, units, vals = rows[0], rows[1], rows[2]
    return {h: (v, u) for h, u, v in zip(hdr, units, vals)}


def main():
    args = sys.argv[1:]
    out = None
    if "--csv" in args:
        i = args.index("--csv")
        out = args[i + 1]
        args = args[:i] + args[i + 2:]
    tabs = [(f, load(f)) for f in args]
    lines = [["metric", "unit"] + [d.get("Kernel Name", ("?",))[0] for _, d in tabs]]
    for k in WANT:
        if any(k in d for _, d in tabs):
            lines.append([k, next((d[k][1] for _, d in tabs if k in d), "")] + [d.get(k, ("",))[0] for _, d in tabs])
    stall = sorted({k for _, d in tabs for k in d if "issue_stalled" in k and k.endswith("per_issue_active.ratio") and "not_issued" not in k})
    for k in stall:
        lines.append(["stall_" + k.split("issue_stalled_")[1].split("_per_issue")[0], "per issue"] + [d.get(k, ("",))[0] for _, d in tabs])
    for ln in lines:
        print("  ".join("%-70s" % ln[0] if i == 0 else "%-14s" % c[:40] for i, c in enumerate(ln)))
    if out:
        csv.writer(open(out, "w")).writerows(lines)


if __name__ == "__main__":
    main()
