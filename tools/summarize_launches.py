"""Summarise an ncu launch list (gpu__time_duration.sum [+ dram bytes]) per kernel: python tools/summarize_launches.py launches.csv [steps]"""
import collections, csv, json, sys

path = sys.argv[1]
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 1
rows = list(csv.reader(open(path)))
hdr = next(i for i, r in enumerate(rows) if "Kernel Name" in r)
H = rows[hdr]
ki, mi, vi, ui, idi = H.index("Kernel Name"), H.index("Metric Name"), H.index("Metric Value"), H.index("Metric Unit"), H.index("ID")
agg = collections.defaultdict(lambda: {"n": 0, "us": 0.0, "rd": 0.0, "wr": 0.0})
scale = {"ns": 1e-3, "us": 1.0, "ms": 1e3, "byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
for r in rows[hdr + 1:]:
    if len(r) <= vi:
        continue
    name = r[ki].split("(")[0].replace("void ", "")
    v = float(r[vi].replace(",", "")) * scale.get(r[ui], 1.0)
    a = agg[name]
    if r[mi] == "gpu__time_duration.sum":
        a["n"] += 1; a["us"] += v
    elif r[mi] == "dram__bytes_read.sum":
        a["rd"] += v
    elif r[mi] == "dram__bytes_write.sum":
        a["wr"] += v
tot = sum(a["us"] for a in agg.values())
print(f"{'kernel':58s} {'launches/step':>13s} {'us/step':>10s} {'share':>6s} {'us/launch':>9s} {'DRAM MB/step':>12s} {'GB/s':>7s}")
for k, a in sorted(agg.items(), key=lambda kv: -kv[1]["us"]):
    mb = (a["rd"] + a["wr"]) / 1e6 / steps
    print(f"{k[:58]:58s} {a['n'] / steps:13.1f} {a['us'] / steps:10.1f} {100 * a['us'] / tot:5.1f}% {a['us'] / a['n']:9.2f} {mb:12.1f} {(a['rd'] + a['wr']) / a['us'] / 1e3 if a['us'] else 0:7.0f}")
n = sum(a["n"] for a in agg.values())
rd = sum(a["rd"] for a in agg.values()); wr = sum(a["wr"] for a in agg.values())
print(f"total: {n / steps:.0f} launches/step, {tot / steps / 1e3:.3f} ms/step of kernel time, DRAM {(rd + wr) / 1e9 / steps:.3f} GB/step (read {rd / 1e9 / steps:.3f}, write {wr / 1e9 / steps:.3f})")
print(json.dumps({"launches_per_step": n / steps, "kernel_ms_per_step": tot / steps / 1e3, "dram_bytes_per_step": (rd + wr) / steps}))
