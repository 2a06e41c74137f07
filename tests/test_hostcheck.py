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


def test_bvh4_matches_brute_force(tmp_path):
    """csrc/pltb_bvh.h + the Traversal state machine of csrc/pltb_scene.cuh, compiled for the host: closest hits (t and primitive,
    bit for bit) and occlusion of random rays equal brute force; the stack never exceeds the bound pltb_scene_create checks."""
    exe = str(tmp_path / "bvh_check")
    subprocess.check_call(["g++", "-O2", "-std=c++17", "-ffp-contract=off", "-I" + os.path.join(ROOT, "mitsuba3_plt_b200", "csrc"),
                           "-I" + os.path.join(ROOT, "include"), "-o", exe, os.path.join(ROOT, "tests", "hostcheck", "bvh_check.cpp")])
    for args in (("3000", "20000", "0"), ("20000", "4000", "1"), ("3", "1000", "0"), ("1", "1000", "0")):
        out = json.loads(subprocess.check_output([exe, *args]).decode().strip().splitlines()[-1])
        assert out["mismatches"] == 0 and out["max_sp"] <= out["stack_need"] <= 64, out
