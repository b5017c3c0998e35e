"""Key counters of every launch in an ncu report as JSON (for profiles/ and for bench.py's roofline object):
   python scripts/ncu_extract.py rep.ncu-rep [rays-per-launch] > profiles/<name>.json"""
import csv, json, subprocess, sys
rep = sys.argv[1]
rays = int(sys.argv[2]) if len(sys.argv) > 2 else None
out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(out.splitlines()))
hdr, units = rows[0], rows[1]
U = dict(zip(hdr, units))
want = {
    "gpu__time_duration.sum": "duration",
    "dram__bytes_read.sum": "dram_read",
    "dram__bytes_write.sum": "dram_write",
    "lts__t_sectors.sum": "l2_sectors",
    "SM_B.TriageCompute.l1tex__t_sectors.sum": "l1_sectors",
    "sm__issue_active.avg.pct_of_peak_sustained_elapsed": "issue_active_pct",
    "smsp__inst_executed.sum": "warp_instructions",
    "smsp__thread_inst_executed_per_inst_executed.ratio": "active_threads_per_warp",
    "smsp__issue_active.avg.pct": "issue_slots_busy_pct",
    "sm__inst_issued.avg.pct_of_peak_sustained_active": "inst_issued_pct_of_peak",
    "sm__warps_active.avg.pct_of_peak_sustained_active": "achieved_occupancy_pct",
    "l1tex__throughput.avg.pct_of_peak_sustained_active": "l1tex_throughput_pct",
    "lts__throughput.avg.pct_of_peak_sustained_elapsed": "l2_throughput_pct",
    "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed": "dram_throughput_pct",
    "launch__registers_per_thread": "registers",
    "launch__grid_size": "grid",
}
scale = {"Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "byte": 1.0, "usecond": 1e-6, "msecond": 1e-3, "second": 1.0, "nsecond": 1e-9, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}
res = []
for r in rows[2:]:
    d = dict(zip(hdr, r))
    e = {"kernel": d.get("Kernel Name", "")[:120]}
    for k, name in want.items():
        if k in d and d[k] != "":
            v = float(d[k].replace(",", ""))
            u = U.get(k, "")
            if u in scale:
                v *= scale[u]
            e[name] = v
    if rays:
        e["rays"] = rays
        if "dram_read" in e:
            e["dram_bytes_per_ray"] = (e["dram_read"] + e["dram_write"]) / rays
        if "l2_sectors" in e:
            e["l2_bytes_per_ray"] = 32 * e["l2_sectors"] / rays
        if "l1_sectors" in e:
            e["l1_sectors_per_ray"] = e["l1_sectors"] / rays
        if "warp_instructions" in e:
            e["warp_instructions_per_ray"] = e["warp_instructions"] / rays
    res.append(e)
print(json.dumps({"report": rep.split("/")[-1], "launches": res}, indent=1))
