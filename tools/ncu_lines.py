#!/usr/bin/env python
"""Attribute the per-SASS-instruction samples of `ncu --page source --csv` to source lines.

usage: python tools/ncu_lines.py <source.csv[.gz]> <nvdisasm -g -c listing> [kernel substring] [launch index] [top N]

The ncu CSV has one block per profiled launch ("Kernel Name" row, header row, one row per instruction whose address is
absolute); the nvdisasm listing gives `//## File "...", line N` markers per instruction offset.  Offsets are matched by
instruction order within the function."""
import collections
import csv
import gzip
import re
import sys


def listing(path):
    """mangled function -> list of (file, line) per instruction, in order"""
    out, cur, loc = {}, None, ("?", 0)
    for ln in open(path, errors="replace"):
        m = re.match(r"\s*\.section\s+\.text\.(\S+?),", ln)
        if m:
            cur = out.setdefault(m.group(1), [])
            continue
        m = re.match(r'\s*//## File "(.*)", line (\d+)', ln)
        if m:
            loc = (m.group(1).split("/")[-1], int(m.group(2)))
            continue
        if cur is not None and re.match(r"\s*/\*[0-9a-f]{4,}\*/", ln):
            cur.append(loc)
    return out


def blocks(path):
    f = gzip.open(path, "rt") if path.endswith(".gz") else open(path)
    name, hdr, rows = None, None, []
    for r in csv.reader(f):
        if r and r[0] == "Kernel Name":
            if name:
                yield name, hdr, rows
            name, hdr, rows = r[1], None, []
        elif r and r[0] == "Address":
            hdr = r
        elif name and hdr and r:
            rows.append(r)
    if name:
        yield name, hdr, rows


def main():
    src, lst = sys.argv[1], sys.argv[2]
    want = sys.argv[3] if len(sys.argv) > 3 else ""
    which = int(sys.argv[4]) if len(sys.argv) > 4 else 0
    top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
    L = listing(lst)
    k = 0
    for name, hdr, rows in blocks(src):
        if want not in name:
            continue
        if k != which:
            k += 1
            continue
        key = name.split("(")[0].split("::")[-1].split("<")[0]
        cand = [f for f in L if key in f and len(L[f]) == len(rows)]
        if not cand:
            cand = [f for f in L if key in f]
            print("# instruction count mismatch", len(rows), [len(L[f]) for f in cand])
        locs = L[cand[0]]
        iS, iI, iT = hdr.index("# Samples"), hdr.index("Instructions Executed"), hdr.index("Thread Instructions Executed")
        agg = collections.defaultdict(lambda: [0, 0, 0])
        tot = [0, 0, 0]
        for n, r in enumerate(rows):
            loc = locs[n] if n < len(locs) else ("?", 0)
            v = (int(r[iS] or 0), int(r[iI] or 0), int(r[iT] or 0))
            for j in range(3):
                agg[loc][j] += v[j]
                tot[j] += v[j]
        print(f"# {name[:100]}: {len(rows)} instructions, samples {tot[0]}, warp inst {tot[1]}, "
              f"avg threads/inst {tot[2] / max(1, tot[1]):.2f}")
        for loc, v in sorted(agg.items(), key=lambda x: -x[1][0])[:top]:
            print(f"{loc[0]:>24s}:{loc[1]:<5d} samples {100 * v[0] / max(1, tot[0]):5.1f}%  inst {100 * v[1] / max(1, tot[1]):5.1f}%  "
                  f"thr/inst {v[2] / max(1, v[1]):5.1f}")
        return
    print("no such kernel/launch")


if __name__ == "__main__":
    main()
