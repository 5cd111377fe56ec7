"""Text summary of an .ncu-rep for profiles/: python tools/ncu_summary.py rep.ncu-rep > profiles/xxx_ncu.txt"""
import csv, io, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(raw)))
hdr, units = rows[0], rows[1]
want = ["Kernel Name", "Grid Size", "Block Size", "gpu__time_duration.sum", "sm__cycles_elapsed.avg", "smsp__cycles_active.avg",
        "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed", "lts__t_sector_hit_rate.pct", "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed", "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed",
        "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "smsp__inst_executed.sum", "l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum"]
print(f"# ncu --set full --clock-control none --import-source on; source: {rep.split('/')[-1]} (cold caches, serialised launches)")
for w in want:
    if w in hdr:
        i = hdr.index(w)
        print(f"{w} [{units[i]}]: " + " | ".join(r[i][:60] for r in rows[2:]))
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(src)))
secs = [i for i, r in enumerate(rows) if r and r[0] == "Kernel Name"] + [len(rows)]
if len(secs) > 1:
    s, e = secs[0], secs[1]
    h = rows[s + 1]
    iS, iSrc, iE = h.index("# Samples"), h.index("Source"), h.index("Instructions Executed")
    stall = [i for i, x in enumerate(h) if x.startswith("stall_") and "Not Issued" not in x]
    data = [r for r in rows[s + 2:e] if len(r) > max(stall)]
    tot = sum(int(r[iS]) for r in data)
    agg = {h[i]: sum(int(r[i]) for r in data) for i in stall}
    print(f"\n# first launch, warp-state samples: {tot}; by reason: " + ", ".join(f"{k}={v}" for k, v in sorted(agg.items(), key=lambda kv: -kv[1])[:7]))
    print("# top instructions by samples (index, samples, executed, SASS, top two stall reasons)")
    for idx, r in sorted(sorted(enumerate(data), key=lambda kv: -int(kv[1][iS]))[:14]):
        st = sorted(((int(r[i]), h[i]) for i in stall), reverse=True)[:2]
        print(f"{idx:5d} {int(r[iS]):6d} {int(r[iE]):9d}  {r[iSrc].strip()[:60]:60s} {st}")
