"""Active lanes per executed instruction by source region of the quadruped kernel: where does SIMT divergence cost?

    ncu -i rep.ncu-rep --page source --csv --print-source sass > sass.csv
    cuobjdump -xelf all ps_variant_legged_sparse.o ; nvdisasm -g -c *.cubin > kern.sass
    python tools/ncu_lanes_by_region.py sass.csv kern.sass

Regions of ps_coop_solver.cuh are found by the line numbers of its functions in the CURRENT tree, so run it on a
capture of the current build.
"""
import collections
import csv
import os
import re
import sys

sass_csv, disasm = sys.argv[1], sys.argv[2]
solver = os.path.join(os.path.dirname(os.path.abspath(__file__)), "..", "mujoco_mpc_b200", "csrc", "ps_coop_solver.cuh")
marks = []          # (first line, region name), ascending
names = {"co_update_rows": "solver: update_rows", "co_update_cost": "solver: update_cost / J'f", "co_cone_block_single": "solver: cone block",
         "co_newton_gradient": "solver: gradient", "co_newton_direction": "solver: Hessian assembly + Cholesky + solves",
         "co_line_eval": "solver: line eval", "co_line_search": "solver: line search", "co_constraint_stage": "solver: Newton loop"}
for n, line in enumerate(open(solver), 1):
    m = re.match(r"PS_\w+ \w+ (co_\w+)\(", line)
    if m and m.group(1) in names:
        marks.append((n - 2, names[m.group(1)]))


def region(f, l):
    if f != "ps_coop_solver.cuh":
        return f
    name = "solver: (head)"
    for first, nm in marks:
        if l >= first:
            name = nm
    return name


rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]; ci = {h: k for k, h in enumerate(hdr)}
data = rows[2:]
base = int(data[0][ci["Address"]], 16)
loc = ("?", 0); where = {}; sec = -1
for line in open(disasm):
    if line.startswith(".text."):
        sec += 1; continue
    if sec != 0:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        loc = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*);", line)
    if m:
        where[int(m.group(1), 16)] = loc
wi = collections.Counter(); ti = collections.Counter(); sm = collections.Counter()
for r in data:
    f, l = where.get(int(r[ci["Address"]], 16) - base, ("?", 0))
    k = region(f, l)
    wi[k] += int(r[ci["Instructions Executed"]]); ti[k] += int(r[ci["Thread Instructions Executed"]]); sm[k] += int(r[ci["# Samples"]])
W, T, S = sum(wi.values()), sum(ti.values()), sum(sm.values())
print(f"{'region':46s} warp-instr %   lanes of 32   stall samples %")
for k, v in wi.most_common(18):
    print(f"{k:46s} {100 * v / W:10.2f}   {ti[k] / max(v, 1):10.2f}   {100 * sm[k] / max(S, 1):12.2f}")
print(f"{'all':46s} {W:10.4g}   {T / W:10.2f}")
