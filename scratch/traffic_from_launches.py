"""ncu launch list (--metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum --csv) + the JSON line of
the same bench command run without ncu -> DRAM bytes per unit of work for the dominant kernels.
usage: traffic_from_launches.py launches.csv plain.json"""
import csv, json, sys
from collections import defaultdict

rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
hdr = rows[0]
ki, mi, vi, ui = hdr.index("Kernel Name"), hdr.index("Metric Name"), hdr.index("Metric Value"), hdr.index("Metric Unit")
scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-6, "us": 1e-3, "ms": 1, "s": 1e3, "usecond": 1e-3, "msecond": 1, "nsecond": 1e-6, "second": 1e3}
agg = defaultdict(lambda: [0.0, 0.0, 0])          # kernel -> [ms, dram bytes, launches]
for r in rows[1:]:
    name = r[ki].split("<")[0].split("(")[0].replace("void ", "").replace("ptb::", "").strip()
    v = float(r[vi].replace(",", "")) * scale[r[ui]]
    if r[mi] == "gpu__time_duration.sum":
        agg[name][0] += v; agg[name][2] += 1
    else:
        agg[name][1] += v
line = [json.loads(l) for l in open(sys.argv[2]) if l.startswith("{")][-1]
frames = line["steps"] + line["warmup"] + 2          # + the e2e leg's 1 + steps... printed for orientation only
tot = sum(a[0] for a in agg.values())
print(f"{'kernel':28s} {'launches':>8s} {'ms':>10s} {'share':>7s} {'DRAM GB':>9s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1][0]):
    print(f"{k:28s} {a[2]:8d} {a[0]:10.3f} {100 * a[0] / tot:6.1f}% {a[1] / 1e9:9.3f}")
print(json.dumps({k: {"ms": a[0], "dram_bytes": a[1], "launches": a[2]} for k, a in agg.items()}))
