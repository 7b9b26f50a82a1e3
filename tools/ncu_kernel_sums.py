"""Sum gpu__time_duration per kernel name from an `ncu --csv --metrics gpu__time_duration.sum` log, optionally split
into groups at every launch of a marker kernel (e.g. wf_init_kernel)."""
import csv, sys, collections, re
path, marker = sys.argv[1], (sys.argv[2] if len(sys.argv) > 2 else None)
rows = []
with open(path, newline="") as f:
    lines = [l for l in f if l.startswith('"')]
for r in csv.DictReader(lines):
    if r.get("Metric Name") != "gpu__time_duration.sum":
        continue
    v = float(r["Metric Value"].replace(",", ""))
    unit = r["Metric Unit"]
    v *= {"ns": 1e-6, "us": 1e-3, "ms": 1.0, "s": 1e3}.get(unit, 1e-6)
    rows.append((re.sub(r"\(.*", "", r["Kernel Name"]), v))
group, sums = 0, collections.OrderedDict()
for name, ms in rows:
    if marker and marker in name:
        group += 1
    k = (group, name)
    n, t = sums.get(k, (0, 0.0))
    sums[k] = (n + 1, t + ms)
for (g, name), (n, t) in sums.items():
    print(f"{g:3d} {name[:70]:70s} {n:6d} {t:10.3f} ms")
