"""Summarise ncu reports for profiles/: python tools/ncu_summary.py out.txt rep1.ncu-rep [rep2 ...] [--traffic out.json]

Per kernel launch in each report: duration, DRAM bytes read / written, L2 / L1 sector counts and hit rates, issue and
pipe utilisation, occupancy, registers, instruction count and the top warp-stall reasons (from `ncu --page raw --csv`)."""
import csv
import io
import json
import subprocess
import sys

KEYS = [
    "gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sectors.sum", "lts__t_sector_hit_rate.pct",
    "l1tex__t_sector_hit_rate.pct", "l1tex__t_sectors_pipe_lsu_mem_global_op_ld.sum",
    "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
    "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_subpipe_tcgen05_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
    "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_eligible.avg.per_cycle_active",
    "smsp__inst_executed.sum", "launch__registers_per_thread", "launch__grid_size", "launch__block_size",
    "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum",
]


def rows_of(rep):
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    return hdr, rows[1], [r for r in rows[2:] if len(r) == len(hdr)]


MULT = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12}


def num(x):
    try:
        return float(x.replace(",", ""))
    except ValueError:
        return None


def main():
    args = sys.argv[1:]
    traffic_out = None
    if "--traffic" in args:
        i = args.index("--traffic")
        traffic_out = args[i + 1]
        args = args[:i] + args[i + 2:]
    out_path, reps = args[0], args[1:]
    lines, traffic = [], {}
    for rep in reps:
        hdr, units, rows = rows_of(rep)
        col = {h: i for i, h in enumerate(hdr)}
        lines.append("== %s" % rep)
        for r in rows:
            name = r[col["Kernel Name"]]
            lines.append("  Kernel Name  %s" % name[:150])
            for k in KEYS:
                if k in col:
                    lines.append("    %-82s %s %s" % (k, r[col[k]], units[col[k]]))
            st = [(num(r[i]), h) for h, i in col.items() if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")]
            st = sorted([s for s in st if s[0] is not None], reverse=True)[:5]
            lines.append("    top stalls (warps per issue): " + ", ".join("%s %.2f" % (h[len("smsp__average_warps_issue_stalled_"):-len("_per_issue_active.ratio")], v) for v, h in st))
            rd, wr = num(r[col["dram__bytes_read.sum"]]), num(r[col["dram__bytes_write.sum"]])
            if rd is not None and wr is not None:
                rd *= MULT.get(units[col["dram__bytes_read.sum"]], 1.0)
                wr *= MULT.get(units[col["dram__bytes_write.sum"]], 1.0)
                traffic.setdefault(name.split("(")[0].strip(), []).append((rd, wr))
        lines.append("")
    open(out_path, "w").write("\n".join(lines) + "\n")
    if traffic_out:
        json.dump({k: {"launches": len(v), "dram_read": [a for a, _ in v], "dram_write": [b for _, b in v]} for k, v in traffic.items()},
                  open(traffic_out, "w"), indent=1)
    print("wrote", out_path)


if __name__ == "__main__":
    main()
