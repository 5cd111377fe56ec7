"""Top stalled instructions of one kernel in an `ncu --page source --csv` export:
   ncu -i rep --page source --csv > src.csv; python tools/ncu_src_top.py src.csv [section-index] [n]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
secs = []
for i, r in enumerate(rows):
    if r and r[0] == "Kernel Name":
        secs.append(i)
secs.append(len(rows))
k = int(sys.argv[2]) if len(sys.argv) > 2 else 0
n = int(sys.argv[3]) if len(sys.argv) > 3 else 40
print("sections:", [(j, rows[s][1][:70]) for j, s in enumerate(secs[:-1])])
s, e = secs[k], secs[k + 1]
hdr = rows[s + 1]
iS, iSrc, iE = hdr.index("# Samples"), hdr.index("Source"), hdr.index("Instructions Executed")
stall = [i for i, h in enumerate(hdr) if h.startswith("stall_") and "Not Issued" not in h]
data = [r for r in rows[s + 2:e] if len(r) > max(stall)]
tot = sum(int(r[iS]) for r in data)
print(rows[s][1], "samples", tot, "SASS lines", len(data), "warp instr executed", sum(int(r[iE]) for r in data))
agg = {hdr[i]: sum(int(r[i]) for r in data) for i in stall}
print(", ".join(f"{k}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:8]))
top = sorted(enumerate(data), key=lambda kv: -int(kv[1][iS]))[:n]
for idx, r in sorted(top):
    st = sorted(((int(r[i]), hdr[i]) for i in stall), reverse=True)[:2]
    print(f"{idx:5d} {int(r[iS]):6d} {int(r[iE]):8d}  {r[iSrc].strip()[:64]:64s} {st}")
