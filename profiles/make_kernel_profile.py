#!/usr/bin/env python
"""profiles/r2_counters.csv (ncu per-launch counters of one `bench.py --steps 1 --warmup 1` run, see capture_r2.sh) +
the bench line of the same command -> profiles/r2_kernel_profile.json, the file bench.py's `roofline` reads:
per kernel: launches, ms, DRAM / L2 bytes per launch, warp and thread instructions, active threads per instruction;
for the closest-hit traversal also thread instructions per ray (rays from the bench line's own counters).
usage: make_kernel_profile.py <r2_counters.csv> <r2_plain.json> <out.json>"""
import collections
import csv
import json
import re
import sys


def short(name):
    name = re.sub(r"\(.*", "", name.replace("mtsb::", "").replace("void ", ""))
    name = name.replace("(bool)0", "0").replace("(bool)1", "1")
    # triangle-only instances (second template argument Shapes = 0) keep their round-2 names: k_trace<0>, k_trace<1>, k_shadow
    return name.replace("k_trace<0, 0>", "k_trace<0>").replace("k_trace<1, 0>", "k_trace<1>").replace("k_shadow<0>", "k_shadow")


rows = [r for r in csv.reader(l for l in open(sys.argv[1]) if l.startswith('"'))]
h = rows[0]
iK, iM, iV, iI = h.index("Kernel Name"), h.index("Metric Name"), h.index("Metric Value"), h.index("ID")
per = collections.OrderedDict()
for r in rows[1:]:
    per.setdefault((r[iI], short(r[iK])), {})[r[iM]] = float(r[iV].replace(",", ""))
# a bench run with --steps 1 --warmup 1 holds 2 render steps (+ the e2e step): keep every launch, report per-launch means
agg = collections.OrderedDict()
for (_, k), m in per.items():
    a = agg.setdefault(k, collections.Counter())
    a["launches"] += 1
    for key, val in m.items():
        a[key] += val
bench = json.loads(open(sys.argv[2]).read().strip().splitlines()[-1])
out = {}
for k, a in agg.items():
    n = a["launches"]
    d = {"launches": int(n), "ms_per_launch": a["gpu__time_duration.sum"] / n / 1e6,
         "dram_bytes_per_launch": (a["dram__bytes_read.sum"] + a["dram__bytes_write.sum"]) / n,
         "l2_bytes_per_launch": a["lts__t_bytes.sum"] / n, "warp_inst_per_launch": a["smsp__inst_executed.sum"] / n,
         "thread_inst_per_launch": a["smsp__thread_inst_executed.sum"] / n,
         "threads_per_inst": a["smsp__thread_inst_executed.sum"] / max(1.0, a["smsp__inst_executed.sum"]),
         "issue_slots_busy_pct": a["smsp__issue_active.avg.pct_of_peak_sustained_active"] / n}
    out[k] = d
if "k_trace<0>" in out:
    # rays per closest-hit launch of this workload = the bench line's own count (rays per sample x samples per step / launches per step)
    rays_per_step = bench["extra"]["rays_per_sample"] * bench["config"]["samples_per_step"]
    launches_per_step = 6
    t = out["k_trace<0>"]
    t["rays_per_launch"] = rays_per_step / launches_per_step
    t["thread_inst_per_ray"] = t["thread_inst_per_launch"] / t["rays_per_launch"]
    t["algorithmic_bytes_per_launch"] = 48 * t["rays_per_launch"]
out["_source"] = {"command": "python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-side-metrics --strong-spp 0", "bench_value": bench["value"]}
json.dump(out, open(sys.argv[3], "w"), indent=1)
print(json.dumps({k: v for k, v in out.items() if k.startswith("k_trace")}, indent=1))
