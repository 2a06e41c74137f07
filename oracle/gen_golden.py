#!/usr/bin/env python
"""Regenerates tests/golden/ref_kat.json and ref_kat_pol.json from the reference's own headers (container only: needs /root/reference).
    python oracle/gen_golden.py [/root/reference]
Builds oracle/_ref/ref_kat with oracle/ref_overlay.sh, runs it, and stores its JSON output (non-finite values spelled the way
Python's json module reads them)."""
import json, os, re, subprocess, sys
HERE = os.path.dirname(os.path.abspath(__file__))
ref = sys.argv[1] if len(sys.argv) > 1 else "/root/reference"
subprocess.check_call(["sh", os.path.join(HERE, "ref_overlay.sh"), ref, "_ref"])
for exe, out_name in (("ref_kat", "ref_kat.json"), ("ref_kat_pol", "ref_kat_pol.json")):
    raw = subprocess.check_output([os.path.join(HERE, "_ref", exe)]).decode()
    raw = re.sub(r"-?nan", "NaN", raw)
    raw = re.sub(r"(?<![\w.])-inf", "-Infinity", raw)
    raw = re.sub(r"(?<![\w.-])inf", "Infinity", raw)
    data = json.loads(raw)
    data["_generator"] = (f"oracle/{exe}.cpp compiled against the reference headers (ZenithGD/mitsuba3-plt @ bc454f3d) by "
                          "oracle/gen_golden.py")
    out = os.path.join(HERE, "..", "tests", "golden", out_name)
    with open(out, "w") as f:
        json.dump(data, f, separators=(",", ":"))
    print({k: (len(v) if isinstance(v, list) else v) for k, v in data.items()}, os.path.getsize(out))
