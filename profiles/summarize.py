#!/usr/bin/env python
"""Turn ncu outputs into the small text summaries committed under profiles/.

  summarize.py launches <launch_list.csv>            per-kernel share of the step (gpu__time_duration.sum pass)
  summarize.py raw <report.ncu-rep>                  key metrics of every captured launch (ncu --set full)
  summarize.py hot <report.ncu-rep> <launch#> [n]    top-n stall-sample SASS lines of one captured launch
  summarize.py bysource <report.ncu-rep> <kernel regex> [n]   warp instructions of the first matching launch per CUDA source file,
                                                     and its top-n source lines (needs -lineinfo and --import-source on)
"""
import collections
import csv
import io
import re
import subprocess
import sys

KEYS = [
    ("gpu__time_duration.sum", "time"), ("launch__grid_size", "grid"), ("launch__block_size", "block"),
    ("launch__registers_per_thread", "regs"), ("sm__warps_active.avg.pct_of_peak_sustained_active", "occupancy %"),
    ("smsp__issue_active.avg.pct_of_peak_sustained_active", "issue slots busy %"),
    ("smsp__thread_inst_executed_per_inst_executed.ratio", "threads/inst"), ("smsp__inst_executed.sum", "warp insts"),
    ("dram__bytes_read.sum", "dram read"), ("dram__bytes_write.sum", "dram write"),
    ("l1tex__t_sector_hit_rate.pct", "L1 hit %"), ("lts__t_sector_hit_rate.pct", "L2 hit %"),
    ("sm__throughput.avg.pct_of_peak_sustained_elapsed", "SM throughput %"),
    ("gpu__compute_memory_throughput.avg.pct_of_peak_sustained_elapsed", "mem throughput %"),
]


def short(name):
    name = re.sub(r"<cub::policy_hub.*", "", re.sub(r"\(.*", "", name.replace("mtsb::", "").replace("void ", "")))
    return name.replace("(bool)0", "closest").replace("(bool)1", "any")


def launches(path):
    rows = [r for r in csv.reader(l for l in open(path) if l.startswith('"'))]
    h = rows[0]
    iN, iV, iG = h.index("Kernel Name"), h.index("Metric Value"), h.index("Grid Size")
    tot = collections.OrderedDict()
    for r in rows[1:]:
        k = short(r[iN])
        t = tot.setdefault(k, [0, 0.0])
        t[0] += 1
        t[1] += float(r[iV].replace(",", "")) / 1e6
    s = sum(v[1] for v in tot.values())
    print(f"{'kernel':44s} {'launches':>8s} {'ms':>10s} {'share':>7s}")
    for k, (n, ms) in sorted(tot.items(), key=lambda kv: -kv[1][1]):
        print(f"{k:44s} {n:8d} {ms:10.3f} {100 * ms / s:6.1f}%")
    print(f"{'total':44s} {sum(v[0] for v in tot.values()):8d} {s:10.3f}")


def ncu_csv(rep, page, extra=()):
    out = subprocess.run(["ncu", "-i", rep, "--page", page, "--csv", *extra], capture_output=True, text=True).stdout
    return list(csv.reader(io.StringIO(out)))


def raw(rep):
    rows = ncu_csv(rep, "raw")
    h, u = rows[0], rows[1]
    iN = h.index("Kernel Name")
    for n, r in enumerate(rows[2:]):
        print(f"--- launch {n}: {short(r[iN])}")
        for key, label in KEYS:
            if key in h:
                i = h.index(key)
                print(f"    {label:22s} {r[i]:>18s} {u[i]}")


def hot(rep, which, n=25):
    rows = ncu_csv(rep, "source", ["--print-source", "sass"])
    blocks, cur = [], None
    for r in rows:
        if r and r[0] == "Kernel Name":
            cur = {"name": r[1], "rows": []}
            blocks.append(cur)
        elif r and r[0] == "Address":
            cur["hdr"] = r
        elif cur is not None and r:
            cur["rows"].append(r)
    blocks = blocks[::2] if len(blocks) > 1 and blocks[0]["name"] == blocks[1]["name"] and len(blocks) % 2 == 0 else blocks
    b = blocks[which]
    h = b["hdr"]
    iS, iE, iT = h.index("# Samples"), h.index("Instructions Executed"), h.index("Avg. Threads Executed")
    ts = sum(int(r[iS]) for r in b["rows"]) or 1
    te = sum(int(r[iE]) for r in b["rows"]) or 1
    print(f"{short(b['name'])}: {len(b['rows'])} SASS lines, {te} warp insts, {ts} stall samples")
    top = sorted(range(len(b["rows"])), key=lambda i: -int(b["rows"][i][iS]))[:n]
    print(f"{'line':>5s} {'samples':>8s} {'%':>5s} {'execs':>10s} {'thr':>4s}  sass")
    for i in sorted(top):
        r = b["rows"][i]
        print(f"{i:5d} {r[iS]:>8s} {100 * int(r[iS]) / ts:5.1f} {r[iE]:>10s} {r[iT]:>4s}  {r[1].strip()[:70]}")


def bysource(rep, kernel, n=30):
    rows = ncu_csv(rep, "source", ["--print-source", "cuda,sass", "--kernel-name", "regex:" + kernel])
    secs, cur, fn, path = [], None, None, None
    for r in rows:
        if not r:
            continue
        if r[0] == "File Path":
            path = r[1]
        elif r[0] == "Function Name":
            fn = r[1]
        elif r[0] == "Line No":
            cur = {"file": path, "fn": fn, "hdr": r, "rows": []}
            secs.append(cur)
        elif cur is not None and r[0] != "" and len(r) > 2 and r[2] == "-":      # a CUDA source line (SASS rows have an address there)
            cur["rows"].append(r)
    if not secs:
        print("no source-level data for", kernel)
        return
    first_fn, seen, lines, per_file, tot = secs[0]["fn"], set(), [], collections.Counter(), 0
    for sec in secs:                       # the report repeats the pages for every captured launch: keep the first one's
        if sec["fn"] != first_fn or sec["file"] in seen:
            continue
        seen.add(sec["file"])
        h = sec["hdr"]
        iE, iT = h.index("Instructions Executed"), h.index("Thread Instructions Executed")
        for r in sec["rows"]:
            try:
                e, t = int(r[iE]), int(r[iT])
            except ValueError:
                continue
            f = sec["file"].split("/")[-1]
            per_file[f] += e
            tot += e
            lines.append((e, t, f, r[0], r[1].strip()[:100]))
    name = short(first_fn.replace("(bool)0", "0").replace("(bool)1", "1"))
    print(f"{name}: warp instructions per CUDA source file (inlined code is charged to the file it was written in)")
    for f, c in per_file.most_common():
        print(f"  {f:28s} {100 * c / max(tot, 1):5.1f} %")
    print(f"top {n} source lines (share of warp instructions, active threads per instruction)")
    for e, t, f, ln, src in sorted(lines, reverse=True)[:n]:
        print(f"  {100 * e / max(tot, 1):5.2f} %  {t / max(e, 1):5.1f}  {f}:{ln}  {src}")


if __name__ == "__main__":
    cmd = sys.argv[1]
    if cmd == "launches":
        launches(sys.argv[2])
    elif cmd == "raw":
        raw(sys.argv[2])
    elif cmd == "bysource":
        bysource(sys.argv[2], sys.argv[3], int(sys.argv[4]) if len(sys.argv) > 4 else 30)
    elif cmd == "hot":
        hot(sys.argv[2], int(sys.argv[3]), int(sys.argv[4]) if len(sys.argv) > 4 else 25)
