"""Markdown summary of one `ncu --set full` capture (raw CSV export) for profiles/.
   python tools/ncu_summary.py raw.csv "<title>" <sample_steps> [sass.csv disasm.sass section] > profiles/xyz.md"""
import csv
import subprocess
import sys

raw, title, steps = sys.argv[1], sys.argv[2], float(sys.argv[3])
rows = list(csv.reader(open(raw))); hdr, units, vals = rows[0], rows[1], rows[2]


def g(k, default=None):
    return (float(vals[hdr.index(k)]), units[hdr.index(k)]) if k in hdr else (default, "")


print(f"# {title}\n")
print(f"Kernel: `{vals[hdr.index('Kernel Name')][:120]}`\n")
keys = ["gpu__time_duration.sum", "launch__grid_size", "launch__block_size", "launch__registers_per_thread",
        "launch__shared_mem_per_block_dynamic", "launch__shared_mem_per_block_static", "launch__occupancy_limit_registers",
        "launch__occupancy_limit_shared_mem", "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__warps_active.avg.per_cycle_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "smsp__thread_inst_executed_per_inst_executed.ratio",
        "sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "dram__throughput.avg.pct_of_peak_sustained_elapsed", "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "sass__inst_executed_local_loads", "sass__inst_executed_local_stores", "sass__inst_executed_shared_loads", "sass__inst_executed_shared_stores",
        "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
print("| metric | value | unit |\n|---|---|---|")
for k in keys:
    v, u = g(k)
    if v is not None:
        print(f"| {k} | {v:.6g} | {u} |")
print("\nWarp stall reasons (cycles per issued instruction, `smsp__average_warps_issue_stalled_*_per_issue_active`):\n")
st = []
for i, h in enumerate(hdr):
    if "issue_stalled" in h and h.endswith("per_issue_active.ratio") and "not_issued" not in h:
        st.append((float(vals[i] or 0), h.split("issue_stalled_")[1].replace("_per_issue_active.ratio", "")))
print("| reason | cycles |\n|---|---|")
for v, n in sorted(st, reverse=True)[:8]:
    print(f"| {n} | {v:.3f} |")
out = subprocess.run([sys.executable, __file__.replace("ncu_summary", "ncu_flops"), raw, str(steps)], capture_output=True, text=True).stdout
print(f"\nPer sample-step ({steps:.0f} sample-steps in this launch): `{out.strip()}`\n")
if len(sys.argv) > 6:
    out = subprocess.run([sys.executable, __file__.replace("ncu_summary", "ncu_hotlines"), sys.argv[4], sys.argv[5], sys.argv[6]], capture_output=True, text=True).stdout
    print("Executed warp-instructions / stall samples by source (tools/ncu_hotlines.py):\n\n```\n" + "\n".join(out.splitlines()[:40]) + "\n```")
