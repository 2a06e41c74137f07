#!/usr/bin/env python
"""Per-kernel summary of an `ncu --set full ... --page raw --csv` export of one wavefront pass.

usage: python tools/ncu_traffic.py <raw.csv[.gz]> <out.txt> [<traffic.json>]

Writes a table (one row per kernel of csrc/pltb_kernels.cu: launches, time, share of the pass, issue-slot / FMA-pipe / XU(MUFU)-pipe
utilisation, DRAM bytes and GB/s, L1/L2 hit rates, active lanes) and, optionally, profiles/ncu_traffic.json =
{kernel: {"dram_bytes_per_launch": ..}} which bench.py reports as `roofline.traffic`."""
import collections, csv, gzip, json, re, sys

# material classes are BSDF plugin types (PLTB_BSDF_*): 5 = roughgrating, everything else is reported as "other" (as bench.py does)
NAMES = [(r"k_raygen", "k_raygen"), (r"k_shade(_pol)?<\(int\)5>|k_shade(_pol)?<5>", "k_shade<grating>"),
         (r"k_shade(_pol)?<", "k_shade<other>"), (r"k_trace<0>|k_trace<\(bool\)0>|k_trace<false>", "k_trace"),
         (r"k_trace<1>|k_trace<\(bool\)1>|k_trace<true>", "k_trace<shadow>"), (r"k_classify", "k_classify"),
         (r"k_nee_light", "k_nee_light"), (r"k_nee_post(_pol)?<\(int\)5|k_nee_post(_pol)?<5", "k_nee_post<grating>"),
         (r"k_nee_post(_pol)?<", "k_nee_post<other>"), (r"k_nee_scan", "k_nee_scan"), (r"k_nee_eval", "k_nee_eval"),
         (r"k_gather", "k_gather"), (r"k_splat", "k_splat"), (r"k_develop", "k_develop")]
COLS = {"t": "gpu__time_duration.sum", "rd": "dram__bytes_read.sum", "wr": "dram__bytes_write.sum",
        "issue": "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "fma": "sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active",
        "xu": "sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active",
        "alu": "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active",
        "l1": "l1tex__t_sector_hit_rate.pct", "l2": "lts__t_sector_hit_rate.pct",
        "lanes": "smsp__thread_inst_executed_per_inst_executed.ratio", "occ": "sm__warps_active.avg.pct_of_peak_sustained_active",
        "regs": "launch__registers_per_thread"}
SCALE = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1, "ms": 1e3, "s": 1e6}


def main():
    src, out = sys.argv[1], sys.argv[2]
    f = gzip.open(src, "rt") if src.endswith(".gz") else open(src)
    r = csv.reader(f)
    hdr, units = next(r), next(r)
    ix = {k: hdr.index(v) for k, v in COLS.items() if v in hdr}
    kn = hdr.index("Kernel Name")
    agg = collections.OrderedDict()
    for row in r:
        name = next((n for pat, n in NAMES if re.search(pat, row[kn])), None)
        if name is None:
            continue
        def val(k):
            if k not in ix or row[ix[k]] in ("", "n/a"):
                return 0.0
            return float(row[ix[k]].replace(",", "")) * SCALE.get(units[ix[k]], 1)
        a = agg.setdefault(name, collections.defaultdict(float))
        t = val("t")
        a["n"] += 1; a["t"] += t; a["rd"] += val("rd"); a["wr"] += val("wr")
        for k in ("issue", "fma", "xu", "alu", "l1", "l2", "lanes", "occ"):
            a[k] += val(k) * t          # time-weighted
        a["regs"] = val("regs")
    T = sum(a["t"] for a in agg.values())
    lines = [f"# {src}: one wavefront pass; time-weighted averages over the launches of each kernel; total {T:.1f} us",
             f"{'kernel':22s} {'n':>3s} {'us':>8s} {'share':>6s} {'issue%':>7s} {'fma%':>6s} {'alu%':>6s} {'xu%':>6s} {'lanes':>6s} "
             f"{'occ%':>5s} {'regs':>5s} {'dramMB':>8s} {'GB/s':>7s} {'L1hit%':>7s} {'L2hit%':>7s}"]
    traffic = {}
    for name, a in agg.items():
        t = max(a["t"], 1e-9)
        mb = (a["rd"] + a["wr"]) / 1e6
        lines.append(f"{name:22s} {int(a['n']):3d} {a['t']:8.1f} {100 * a['t'] / T:5.1f}% {a['issue'] / t:7.1f} {a['fma'] / t:6.1f} "
                     f"{a['alu'] / t:6.1f} {a['xu'] / t:6.1f} {a['lanes'] / t:6.1f} {a['occ'] / t:5.1f} {int(a['regs']):5d} {mb:8.1f} "
                     f"{mb / t * 1e3:7.1f} {a['l1'] / t:7.1f} {a['l2'] / t:7.1f}")
        traffic[name] = {"dram_bytes_per_launch": round((a["rd"] + a["wr"]) / a["n"]), "launches_in_capture": int(a["n"]),
                         "avg_launch_us": round(a["t"] / a["n"], 2), "issue_slots_busy_pct": round(a["issue"] / t, 1),
                         "active_lanes": round(a["lanes"] / t, 1), "occupancy_pct": round(a["occ"] / t, 1), "registers": int(a["regs"])}
    open(out, "w").write("\n".join(lines) + "\n")
    print("\n".join(lines))
    if len(sys.argv) > 3:
        traffic["_source"] = f"ncu --set full --clock-control none, one pass of cbox-grating 1024x1024 x 8 spp ({src})"
        traffic["_samples_per_pass_in_capture"] = int(sys.argv[4]) if len(sys.argv) > 4 else 1024 * 1024 * 8
        json.dump(traffic, open(sys.argv[3], "w"), indent=1)


if __name__ == "__main__":
    main()
