"""Join an ncu SASS source page (ncu -i x.ncu-rep --page source --csv --kernel-name regex:K --launch-count 1) with
nvdisasm line info of the same cubin (nvdisasm -g -c x.cubin) and print where the warp-stall samples and executed
instructions fall, per source file:line. Usage: ncu_source_lines.py <ncu_source.csv> <nvdisasm.sass> <mangled-substring> [top]"""
import collections
import csv
import re
import sys

csv_path, sass_path, kernel = sys.argv[1:4]
top = int(sys.argv[4]) if len(sys.argv) > 4 else 40

# nvdisasm: ordered list of (offset, file, line) for the kernel's section
lines = open(sass_path, errors="ignore").read().splitlines()
start = next(i for i, l in enumerate(lines) if l.startswith("\t.section\t.text.") and kernel in l)
insts, cur = [], ("?", 0)
for l in lines[start + 1:]:
    if l.startswith("\t.section"):
        break
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split("/")[-1], int(m.group(2)))
        continue
    m = re.match(r"\s*/\*([0-9a-f]{4,})\*/\s+(.*?);", l)
    if m:
        insts.append((int(m.group(1), 16), cur, m.group(2).strip()))

rows = list(csv.reader(open(csv_path)))
hdr = rows[1]
col = {h: i for i, h in enumerate(hdr)}
body = [r for r in rows[2:] if len(r) == len(hdr) and r[0].startswith("0x")]
base = int(body[0][col["Address"]], 16)
by_off = {off: (loc, txt) for off, loc, txt in insts}
agg = collections.defaultdict(lambda: [0, 0, 0, 0])  # samples, no_inst, inst executed, thread inst
total = [0, 0, 0, 0]
unmapped = 0
for r in body:
    off = int(r[col["Address"]], 16) - base
    loc = by_off.get(off, (("?", 0), ""))[0]
    if loc == ("?", 0):
        unmapped += 1
    vals = [int(r[col["# Samples"]] or 0), int(r[col["stall_no_inst"]] or 0), int(r[col["Instructions Executed"]] or 0),
            int(r[col["Thread Instructions Executed"]] or 0)]
    for k in range(4):
        agg[loc][k] += vals[k]
        total[k] += vals[k]
print(f"{len(body)} SASS instructions ({len(body) * 16 / 1024:.0f} KiB), {unmapped} without line info; samples {total[0]}, "
      f"no_inst {total[1]} ({100 * total[1] / max(1, total[0]):.1f} %), warp insts {total[2]}, "
      f"avg active threads {total[3] / max(1, total[2]):.1f}")
files = collections.defaultdict(lambda: [0, 0, 0, 0])
for (f, _), v in agg.items():
    for k in range(4):
        files[f][k] += v[k]
print("-- by file: samples%  no_inst%  inst%  threads/inst")
for f, v in sorted(files.items(), key=lambda kv: -kv[1][0]):
    print(f"{f:28s} {100 * v[0] / total[0]:6.1f} {100 * v[1] / max(1, total[1]):6.1f} {100 * v[2] / total[2]:6.1f} "
          f"{v[3] / max(1, v[2]):6.1f}")
print(f"-- top {top} lines by samples")
for (f, ln), v in sorted(agg.items(), key=lambda kv: -kv[1][0])[:top]:
    print(f"{f}:{ln:<5d} {100 * v[0] / total[0]:6.2f} % samples  no_inst {100 * v[1] / max(1, v[0]):5.1f} %  "
          f"inst {100 * v[2] / total[2]:6.2f} %  thr/inst {v[3] / max(1, v[2]):5.1f}")
