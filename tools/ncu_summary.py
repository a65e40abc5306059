"""profiles/ helper: the handful of numbers of an `ncu --set full` capture that the roofline discussion uses.

    python tools/ncu_summary.py <report.ncu-rep> <moves per launch> <out.csv> [note]

Writes metric,unit,value rows plus derived per-move figures (instructions, DRAM bytes, L2 bytes = lts__t_sectors x 32)."""
import csv, subprocess, sys
rep, moves, out = sys.argv[1], float(sys.argv[2]), sys.argv[3]
note = sys.argv[4] if len(sys.argv) > 4 else ""
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(raw.splitlines()))
h, u, v = rows[0], rows[1], rows[2]
want = ("gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__occupancy_limit_shared_mem", "launch__occupancy_limit_registers",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active",
        "dram__bytes_read.sum", "dram__bytes_write.sum", "lts__t_sectors.sum", "lts__t_sectors_op_read.sum", "lts__t_sectors_op_write.sum",
        "lts__t_sector_hit_rate.pct", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum")
vals = {}
with open(out, "w") as f:
    if note:
        f.write(f"# {note}\n")
    f.write("metric,unit,value\n")
    for a, b, c in zip(h, u, v):
        if a in want or (a.startswith("smsp__average_warps_issue_stalled") and a.endswith("per_issue_active.ratio")):
            f.write(f"{a},{b},{c}\n")
            try:
                vals[a] = (float(c.replace(",", "")), b)
            except ValueError:
                pass
    scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
    dram = sum(vals[k][0] * scale.get(vals[k][1], 1.0) for k in ("dram__bytes_read.sum", "dram__bytes_write.sum") if k in vals)
    l2 = vals.get("lts__t_sectors.sum", (0.0, ""))[0] * 32.0
    inst = vals.get("smsp__inst_executed.sum", (0.0, ""))[0]
    f.write(f"derived.moves_per_launch,,{moves:.0f}\n")
    f.write(f"derived.warp_instructions_per_move,,{inst / moves:.1f}\n")
    f.write(f"derived.dram_bytes_per_launch,byte,{dram:.0f}\n")
    f.write(f"derived.dram_bytes_per_move,byte,{dram / moves:.1f}\n")
    f.write(f"derived.l2_bytes_per_move,byte,{l2 / moves:.1f}\n")
print(open(out).read())
