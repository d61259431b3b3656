"""FP64 flops per sample-step from an `ncu --set full` raw export:
   (dadd + dmul + 2 dfma thread instructions) / sample_steps.
   python tools/ncu_flops.py raw.csv <sample_steps>"""
import csv
import sys
rows = list(csv.reader(open(sys.argv[1]))); hdr, vals = rows[0], rows[2]
steps = float(sys.argv[2])
g = lambda k: float(vals[hdr.index(k)])
cyc = g("smsp__cycles_elapsed.max") if "smsp__cycles_elapsed.max" in hdr else g("sm__cycles_elapsed.max")
cnt = {op: g(f"smsp__sass_thread_inst_executed_op_{op}_pred_on.sum.per_cycle_elapsed") * cyc for op in ("dadd", "dmul", "dfma")}
flops = cnt["dadd"] + cnt["dmul"] + 2 * cnt["dfma"]
inst = g("smsp__inst_executed.sum")
print({"flops_per_step": flops / steps, "dadd": cnt["dadd"] / steps, "dmul": cnt["dmul"] / steps, "dfma": cnt["dfma"] / steps,
       "warp_inst_per_step": inst / steps, "thread_inst_per_step": inst * g("smsp__thread_inst_executed_per_inst_executed.ratio") / steps,
       "dram_bytes_per_step": (g("dram__bytes_read.sum") * {"Tbyte": 1e12, "Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[rows[1][hdr.index("dram__bytes_read.sum")]]
                               + g("dram__bytes_write.sum") * {"Tbyte": 1e12, "Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1}[rows[1][hdr.index("dram__bytes_write.sum")]]) / steps,
       "duration_ms": g("gpu__time_duration.sum") * {"s": 1e3, "ms": 1, "us": 1e-3, "ns": 1e-6}.get(rows[1][hdr.index("gpu__time_duration.sum")], 1)})
