"""ncu report -> small JSON summary for profiles/ (one entry per captured launch).
usage: python scripts/ncu_summary.py gpurun_out/X.ncu-rep profiles/X_ncu_summary.json"""
import csv, io, json, subprocess, sys
KEEP = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum", "smsp__inst_executed.sum",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "launch__grid_size",
        "launch__block_size", "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "dram__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__t_sector_hit_rate.pct", "l1tex__t_sector_hit_rate.pct",
        "sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_active",
        "smsp__issue_active.avg.pct_of_peak_sustained_active", "sm__cycles_elapsed.avg"]
rep, out = sys.argv[1], sys.argv[2]
txt = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(io.StringIO(txt)))
hdr, units = rows[0], rows[1]
ks = []
for r in rows[2:]:
    d = dict(zip(hdr, r))
    e = {"kernel": d.get("Kernel Name", "")}
    for k in KEEP:
        if k in d:
            e[k] = d[k] + " " + units[hdr.index(k)]
    ks.append(e)
json.dump({"source": "ncu --set full --clock-control none --import-source on, " + rep, "kernels": ks}, open(out, "w"), indent=1)
print(len(ks), "launches")
