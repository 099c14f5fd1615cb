#!/usr/bin/env python
"""profiles/traffic.json from an `ncu --page raw --csv` dump of one step's kernels (one launch each):
DRAM bytes read + written per kernel, summed for the update (k_seg_*) and the lookup (k_gather*), plus
duration, registers, warp instructions and issue-active for the summary table.
    ncu -i prof.ncu-rep --page raw --csv > raw.csv;  python tools/ncu_traffic.py raw.csv "<source note>" [out.json]"""
import csv, json, re, sys

rows = list(csv.reader(open(sys.argv[1])))
hdr, units, data = rows[0], rows[1], rows[2:]
ix = {h: i for i, h in enumerate(hdr)}
scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}


def val(r, name):
    i = ix[name]
    try:
        v = float(r[i].replace(",", ""))
    except ValueError:
        return None
    return v * scale.get(units[i], 1.0)


kernels = {}
for r in data:
    name = re.sub(r"^void |<unnamed>::|\(.*$", "", r[ix["Kernel Name"]])
    short = name.split("<")[0]
    k = kernels.setdefault(short, {"launches": 0, "dram_read_bytes": 0.0, "dram_write_bytes": 0.0, "duration_us": 0.0,
                                   "warp_instructions": 0.0, "template": name})
    k["launches"] += 1
    k["dram_read_bytes"] += val(r, "dram__bytes_read.sum") or 0
    k["dram_write_bytes"] += val(r, "dram__bytes_write.sum") or 0
    k["duration_us"] += val(r, "gpu__time_duration.sum") or 0
    k["warp_instructions"] += val(r, "smsp__inst_executed.sum") or 0
    k["registers"] = val(r, "launch__registers_per_thread")
    k["issue_active_pct"] = val(r, "smsp__issue_active.avg.pct_of_peak_sustained_active")
    k["warps_active_pct"] = val(r, "sm__warps_active.avg.pct_of_peak_sustained_active")
upd = sum(k["dram_read_bytes"] + k["dram_write_bytes"] for n, k in kernels.items() if n.startswith("k_seg_"))
lkp = sum(k["dram_read_bytes"] + k["dram_write_bytes"] for n, k in kernels.items() if n.startswith("k_gather"))
out = {"source": sys.argv[2] if len(sys.argv) > 2 else sys.argv[1], "update_dram_bytes_per_step": upd,
       "lookup_dram_bytes_per_step": lkp, "kernels": kernels}
path = sys.argv[3] if len(sys.argv) > 3 else "profiles/traffic.json"
json.dump(out, open(path, "w"), indent=1)
print(f"update {upd / 1e6:.1f} MB, lookup {lkp / 1e6:.1f} MB per step -> {path}")
for n, k in sorted(kernels.items(), key=lambda x: -x[1]["duration_us"]):
    print(f"  {n:22s} x{k['launches']} {k['duration_us']:8.1f} us  dram {(k['dram_read_bytes'] + k['dram_write_bytes']) / 1e6:8.1f} MB  "
          f"regs {k['registers']}  inst {k['warp_instructions'] / 1e6:6.1f} M  issue {k['issue_active_pct']}")
