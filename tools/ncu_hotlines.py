"""Join an ncu SASS-page export with nvdisasm line info: executed warp-instructions and stall samples
per source line / per source file.

  ncu -i rep.ncu-rep --page source --csv --print-source sass > sass.csv
  cuobjdump -xelf all obj.o ; nvdisasm -g -c obj.sm_100a.cubin > kern.sass
  python tools/ncu_hotlines.py sass.csv kern.sass [kernel-section-index]
"""
import collections
import csv
import re
import sys

sass_csv, disasm = sys.argv[1], sys.argv[2]
section = int(sys.argv[3]) if len(sys.argv) > 3 else 0
rows = list(csv.reader(open(sass_csv)))
hdr = rows[1]
ci = {h: k for k, h in enumerate(hdr)}
data = rows[2:]
base = int(data[0][ci["Address"]], 16)
cnt = {}
for r in data:
    off = int(r[ci["Address"]], 16) - base
    cnt[off] = (int(r[ci["Instructions Executed"]]), int(r[ci["# Samples"]]), int(r[ci["Thread Instructions Executed"]]))
# disassembly: pick the section-th .text block
secs = []; cur = None; loc = ("?", 0)
for line in open(disasm):
    if line.startswith(".text."):
        cur = []; secs.append(cur); continue
    if cur is None:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', line)
    if m:
        loc = (m.group(1).split("/")[-1], int(m.group(2))); continue
    m = re.match(r"\s+/\*([0-9a-f]{4,})\*/\s+(.*);", line)
    if m:
        cur.append((int(m.group(1), 16), loc, m.group(2).strip()))
sec = secs[section]
by_line = collections.Counter(); by_file = collections.Counter(); samp_line = collections.Counter(); samp_file = collections.Counter()
tot = 0; tots = 0
for off, loc, txt in sec:
    c = cnt.get(off)
    if not c:
        continue
    by_line[loc] += c[0]; by_file[loc[0]] += c[0]; samp_line[loc] += c[1]; samp_file[loc[0]] += c[1]; tot += c[0]; tots += c[1]
print(f"total warp-instructions {tot}, stall samples {tots}, static instructions {len(sec)}")
print("--- per file: instr%  samples%")
for f, c in by_file.most_common():
    print(f"{f:28s} {100 * c / tot:6.2f}  {100 * samp_file[f] / max(tots, 1):6.2f}")
print("--- top lines by executed instructions")
for (f, l), c in by_line.most_common(45):
    print(f"{f}:{l:<5d} {100 * c / tot:6.2f}%  samples {100 * samp_line[(f, l)] / max(tots, 1):6.2f}%")
print("--- top lines by stall samples")
for (f, l), c in samp_line.most_common(45):
    print(f"{f}:{l:<5d} samples {100 * c / max(tots, 1):6.2f}%  instr {100 * by_line[(f, l)] / tot:6.2f}%")
if len(sys.argv) > 4:
    print("--- requested lines (file:line -> warp-instructions, static instrs)")
    static = collections.Counter()
    for off, loc, txt in sec:
        static[loc] += 1
    for spec in sys.argv[4:]:
        f, l = spec.split(":"); l = int(l)
        print(f"{spec:32s} {by_line[(f, l)]:>14d}  static {static[(f, l)]}")
