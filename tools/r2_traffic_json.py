"""ncu CSV of tools/r2_traffic_cmd.py -> the JSON bench.py reads (profiles/r2_dram_traffic.json)."""
import csv, json, re, sys

rows = [r for r in csv.reader(open(sys.argv[1])) if len(r) > 10]
hdr = rows[0]
col = {h: i for i, h in enumerate(hdr)}
launches = {}
for r in rows[1:]:
    key = (r[col["ID"]], r[col["Kernel Name"]])
    v = float(r[col["Metric Value"]].replace(",", ""))
    unit = r[col["Metric Unit"]]
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "Tbyte": 1e12, "ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3,
             "usecond": 1e-3, "msecond": 1, "nsecond": 1e-6, "second": 1e3}.get(unit, 1)
    launches.setdefault(key, {})[r[col["Metric Name"]]] = v * scale
form = int(re.search(r"form \(\d+, (\d+)\)", open(sys.argv[2]).read()).group(1))
out = {"_note": "dram__bytes_read.sum + dram__bytes_write.sum per launch: ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,"
                "lts__t_bytes.sum,gpu__time_duration.sum --clock-control none over tools/r2_traffic_cmd.py (tools/r2_dram_traffic.sh). "
                "bench.py copies `dram_bytes` into roofline.traffic when its launch is the one measured here."}
for (_, name), m in launches.items():
    rec = {"dram_read_bytes": m["dram__bytes_read.sum"], "dram_write_bytes": m["dram__bytes_write.sum"],
           "dram_bytes": m["dram__bytes_read.sum"] + m["dram__bytes_write.sum"], "l2_bytes": m["lts__t_bytes.sum"],
           "duration_ms_under_ncu": m["gpu__time_duration.sum"], "kernel": name[:120]}
    if "render_sorted" in name and m["gpu__time_duration.sum"] > 50:  # the 256-spp launches (choose_tracer's are 8 spp)
        out["c2"] = dict(rec, form=form, width=1920, height=1080, spp_per_launch=256)
    elif "trace_persistent" in name:
        out["c3"] = dict(rec, rays_per_launch=64 << 20)
print(json.dumps(out, indent=1))
