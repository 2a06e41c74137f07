"""Extract the standard colorimetric tables the path needs (CIE 1931 XYZ 5 nm, CIE D65) from the reference tree
into mitsuba3_plt_b200/data/cie_tables.npz.  These are published CIE data, kept as a data file.
Run in the build container only (needs /root/reference): python tools/gen_tables.py
"""
import re, sys, numpy as np
REF = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"

def floats_after(path, marker, n):
    src = open(path).read()
    i = src.index(marker)
    body = src[src.index("{", i) + 1: src.index("};", i)]
    body = re.sub(r"//.*", "", body)
    vals = [float(x.rstrip("f")) for x in re.findall(r"[-+]?\d*\.?\d+(?:[eE][-+]?\d+)?f?", body)]
    assert len(vals) == n, (len(vals), n)
    return np.array(vals, dtype=np.float32)

xyz = floats_after(f"{REF}/src/core/spectrum.cpp", "cie1931_tbl[MI_CIE_SAMPLES * 3]", 95 * 3).reshape(3, 95)
d65 = floats_after(f"{REF}/include/mitsuba/core/spectrum.h", "d65_table[MI_CIE_SAMPLES]", 95)
np.savez("mitsuba3_plt_b200/data/cie_tables.npz", xyz=xyz, d65=d65)
print(xyz[:, :3], d65[:3], xyz[1].sum())
