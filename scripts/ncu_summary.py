"""Summarise an .ncu-rep (read on the CPU box) into a small markdown table for profiles/."""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
idx = {h: i for i, h in enumerate(hdr)}
cols = [("gpu__time_duration.sum", "time"), ("dram__bytes_read.sum", "DRAM read"), ("dram__bytes_write.sum", "DRAM write"),
        ("gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed", "DRAM %peak"), ("sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active", "tensor pipe %"),
        ("sm__warps_active.avg.pct_of_peak_sustained_active", "warps active %"), ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue active %"),
        ("launch__registers_per_thread", "regs/thread"), ("smsp__inst_executed.sum", "warp instrs"),
        ("l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum", "smem bank conflicts")]
print(f"# ncu summary of `{rep.split('/')[-1]}` (`ncu --set full --clock-control none`, B200)\n")
print("| kernel | " + " | ".join(n for _, n in cols) + " |")
print("|---|" + "---|" * len(cols))
for r in rows[2:]:
    name = r[idx["Kernel Name"]].split("(")[0][:60]
    vals = []
    for k, _ in cols:
        if k in idx:
            try:
                v = float(r[idx[k]])
                vals.append(f"{v:,.2f} {units[idx[k]]}".strip())
            except ValueError:
                vals.append(r[idx[k]])
        else:
            vals.append("n/a")
    print(f"| `{name}` | " + " | ".join(vals) + " |")
