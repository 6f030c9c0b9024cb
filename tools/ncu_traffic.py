#!/usr/bin/env python
"""Per-kernel hardware counters of one frame from an `ncu --set full` capture -> profiles/ncu_traffic.json.

  python tools/ncu_traffic.py <file.ncu-rep> <plain bench log of the same command> <workload> [source label]

The capture holds the launches of ONE frame of `bench.py --workload W --spp S --steps 1 ...`; the plain log is the
JSON line the same command printed without ncu (its rays_per_step are the rays of that frame).  Written per workload:
DRAM bytes per ray of k_extend / k_shadow (dram__bytes_read.sum + dram__bytes_write.sum over the frame's launches ÷
rays), and time-weighted issue-slot %, L1 throughput %, L2 hit % and active lanes per instruction of each kernel —
the explanatory counters bench.py quotes beside the roofline fraction.
"""
import csv
import io
import json
import os
import re
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    rep, plain, workload = sys.argv[1], sys.argv[2], sys.argv[3]
    label = sys.argv[4] if len(sys.argv) > 4 else os.path.basename(rep)
    out = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    hdr = rows[0]
    col = {n: hdr.index(n) for n in hdr}
    line = None
    for l in open(plain):
        if l.startswith("{"):
            line = json.loads(l)
    rays = line["rays_per_step"]
    per = {}
    for r in rows[2:]:
        name = re.sub(r"<.*", "", r[col["Kernel Name"]].split("(")[0].replace("ptb::", "").replace("void ", ""))
        f = lambda k: float(r[col[k]].replace(",", "")) if k in col and r[col[k]] not in ("", "n/a") else 0.0
        unit = lambda k: rows[1][col[k]]
        scale = {"byte": 1.0, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}
        d = per.setdefault(name, {"launches": 0, "ns": 0.0, "dram": 0.0, "issue": 0.0, "l1": 0.0, "l2hit": 0.0, "lanes": 0.0,
                                  "winst": 0.0})
        ns = f("gpu__time_duration.sum") * {"ns": 1.0, "us": 1e3, "ms": 1e6, "usecond": 1e3, "msecond": 1e6, "nsecond": 1.0}.get(unit("gpu__time_duration.sum"), 1.0)
        d["launches"] += 1
        d["ns"] += ns
        d["dram"] += f("dram__bytes_read.sum") * scale.get(unit("dram__bytes_read.sum"), 1.0) + \
            f("dram__bytes_write.sum") * scale.get(unit("dram__bytes_write.sum"), 1.0)
        d["issue"] += ns * f("smsp__issue_active.avg.pct_of_peak_sustained_active")
        d["l1"] += ns * f("l1tex__throughput.avg.pct_of_peak_sustained_elapsed")
        d["l2hit"] += ns * f("lts__t_sector_hit_rate.pct")
        w = f("smsp__inst_executed.sum")
        d["winst"] += w
        d["lanes"] += w * f("smsp__thread_inst_executed_per_inst_executed.ratio")
    res = {"source": label, "frame_rays": rays}
    for k, d in per.items():
        t = max(d["ns"], 1.0)
        res[k + "_ms_under_ncu"] = d["ns"] / 1e6
        res[k + "_launches"] = d["launches"]
        res[k + "_dram_bytes"] = d["dram"]
        res[k + "_issue_pct"] = d["issue"] / t
        res[k + "_l1_throughput_pct"] = d["l1"] / t
        res[k + "_l2_hit_pct"] = d["l2hit"] / t
        res[k + "_active_lanes"] = d["lanes"] / max(d["winst"], 1.0)
    if "k_extend" in per and rays.get("extend"):
        res["k_extend_dram_bytes_per_ray"] = per["k_extend"]["dram"] / rays["extend"]
    if "k_shadow" in per and rays.get("shadow"):
        res["k_shadow_dram_bytes_per_ray"] = per["k_shadow"]["dram"] / rays["shadow"]
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    allv = json.load(open(path)) if os.path.exists(path) else {}
    allv[workload] = res
    json.dump(allv, open(path, "w"), indent=1)
    print(json.dumps(res, indent=1))


if __name__ == "__main__":
    main()
