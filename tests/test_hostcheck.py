"""Host-side checks of header logic in csrc/ (compiled as plain C++; test infrastructure only)."""
import json
import os
import subprocess

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_grating_scan_ranges_match_full_loops(tmp_path):
    """GratingLoop's (lx, ly) index ranges list exactly the lobes the full loops of roughgrating.cpp:845-969 would."""
    exe = str(tmp_path / "scan_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I" + os.path.join(ROOT, "mitsuba3_plt_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "include"), "-o", exe, os.path.join(ROOT, "tests", "hostcheck", "scan_check.cpp")])
    out = json.loads(subprocess.check_output([exe, "300000"]).decode().strip().splitlines()[-1])
    assert out["mismatches"] == 0 and out["with_hits"] > 50000
    assert out["lobe_iters_ranged"] * 3 < out["lobe_iters_full"]
