#!/usr/bin/env python
"""Per-kernel table from `ncu --page raw --csv` of an `--set full` capture: time-weighted means over the launches of each kernel of
issue-slot utilisation, pipe utilisation, active lanes, occupancy, registers, DRAM bytes, cache hit rates and the top warp-stall
reasons (warps stalled per issue-active cycle).  usage: python tools/ncu_table.py <raw.csv[.gz]> [--json out.json]"""
import collections, csv, gzip, json, re, sys

path = sys.argv[1]
f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
r = csv.reader(f)
hdr, units = next(r), next(r)
col = {h: i for i, h in enumerate(hdr)}
M = {"us": "gpu__time_duration.sum", "issue": "sm__issue_active.avg.pct_of_peak_sustained_elapsed",
     "fma": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active", "alu": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
     "xu": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active", "lsu": "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
     "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio", "occ": "sm__warps_active.avg.pct_of_peak_sustained_active",
     "regs": "launch__registers_per_thread", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
     "l1": "l1tex__t_sector_hit_rate.pct", "l2": "lts__t_sector_hit_rate.pct"}
stall_cols = {re.search(r"issue_stalled_(.*)_per_issue_active", h).group(1): i for i, h in enumerate(hdr)
              if h.startswith("smsp__average_warps_issue_stalled_") and h.endswith("_per_issue_active.ratio")}


def num(row, name):
    i = col.get(name)
    if i is None or row[i] in ("", "n/a"):
        return 0.0
    v = float(row[i].replace(",", ""))
    u = units[i]
    if name.startswith("dram__bytes"):
        v *= {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9}.get(u, 1)
    if name == "gpu__time_duration.sum":
        v *= {"ns": 1e-3, "us": 1, "usecond": 1, "ms": 1e3, "msecond": 1e3, "s": 1e6, "second": 1e6, "nsecond": 1e-3}.get(u, 1)
    return v


def short(name):
    m = re.match(r"(?:void )?(?:pltb::)?(k_\w+)(?:<(.*?)>)?", name)
    if not m:
        return name[:24]
    k, t = m.group(1), m.group(2)
    # material class = BSDF plugin type (include/pltb.h): 5 = roughgrating; "--per-class" keeps one row per class
    names = {"0": "null", "1": "diffuse", "2": "delta", "3": "dielectric", "4": "roughconductor", "5": "grating", "6": "principled"}
    if k in ("k_shade", "k_nee_post", "k_shade_pol", "k_nee_post_pol"):
        c = re.sub(r"\D", "", (t or "").split(",")[0])   # first template argument (k_nee_post<class, rgb>)
        if "--per-class" in sys.argv: return f"{k}<{names.get(c, c)}>"
        return f"{k}<grating>" if c == "5" else f"{k}<other>"
    if k == "k_trace": return "k_trace<shadow>" if t.startswith("1") or t.startswith("true") else "k_trace"
    return k


agg = collections.OrderedDict()
for row in r:
    if len(row) < len(hdr):
        continue
    k = short(row[col["Kernel Name"]])
    us = num(row, M["us"])
    a = agg.setdefault(k, collections.defaultdict(float))
    a["n"] += 1; a["us"] += us
    for key in ("issue", "fma", "alu", "xu", "lsu", "lanes", "occ", "l1", "l2"):
        a[key] += us * num(row, M[key])
    a["regs"] = max(a["regs"], num(row, M["regs"]))
    a["dram"] += num(row, M["rd"]) + num(row, M["wr"])
    for s, i in stall_cols.items():
        a["st_" + s] += us * float(row[i].replace(",", "") or 0)
tot = sum(a["us"] for a in agg.values())
print(f"{'kernel':22s} {'n':>3s} {'us':>9s} {'share':>6s} {'issue%':>6s} {'fma%':>5s} {'alu%':>5s} {'xu%':>5s} {'lsu%':>5s} {'lanes':>5s} {'occ%':>5s} "
      f"{'regs':>4s} {'dramMB':>8s} {'GB/s':>7s} {'L1%':>5s} {'L2%':>5s}  top stalls (warps stalled per issue-active cycle)")
out = {}
for k, a in agg.items():
    us = a["us"]
    w = lambda key: a[key] / us if us else 0.0
    st = sorted(((a["st_" + s] / us, s) for s in stall_cols), reverse=True)[:4]
    print(f"{k:22s} {int(a['n']):3d} {us:9.1f} {100 * us / tot:5.1f}% {w('issue'):6.1f} {w('fma'):5.1f} {w('alu'):5.1f} {w('xu'):5.1f} {w('lsu'):5.1f} "
          f"{w('lanes'):5.1f} {w('occ'):5.1f} {int(a['regs']):4d} {a['dram'] / 1e6:8.1f} {a['dram'] / us / 1e3:7.1f} {w('l1'):5.1f} {w('l2'):5.1f}  "
          + "  ".join(f"{s} {v:.2f}" for v, s in st))
    out[k] = {"launches": int(a["n"]), "us": us, "issue_pct": w("issue"), "lanes": w("lanes"), "occupancy_pct": w("occ"),
              "dram_bytes_per_launch": a["dram"] / a["n"], "stalls": {s: v for v, s in st}}
if "--json" in sys.argv:
    json.dump(out, open(sys.argv[sys.argv.index("--json") + 1], "w"), indent=1)
