#!/usr/bin/env python
"""Where do the non-finite radiance samples of the unpolarized spectral `plt` path come from?  

Runs a scene through the CPU oracle with its non-finite trace switched on (PLT_ORACLE_NAN_TRACE, oracle/plt_oracle.cpp
`trace_nonfinite`): every non-finite term of plt_solve_phase is dumped with the factor that is not finite and the state of the
path (per bounce: BSDF type, |wi|, |wo|, sampled lobe, diffract()'s validity mask).  The summary classifies each record by its
root cause and is what DESIGN.md section 5 cites.

    python tools/nan_trace.py cbox-grating --res 256 --spp 16 > profiles/r02_nan_trace_cbox-grating.txt
    python tools/nan_trace.py /root/reference/scenes/disk/disk.xml --res 200x150 --spp 64 --sky
"""
import argparse
import collections
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def classify(rec):
    """root cause of one non-finite term"""
    path = rec["path"]
    invalid = [j for j, b in enumerate(path[:-1]) if not b["lobe_valid"]]
    nonunit_in = abs(path[-1]["wi_norm"] - 1.0) > 1e-3
    if invalid and nonunit_in and rec["first_nonfinite_factor"] == "bsdf_val":
        return ("invalid lobe at bounce %d -> non-unit ray direction -> |wi| != 1 at this vertex -> fresnel_conductor(cos > 1) "
                "overflows (fresnel.h:93-124: term_3 + term_4 cancels) -> wbsdf_eval = inf * 0" % invalid[0])
    if invalid and rec["first_nonfinite_factor"] == "replay_alpha":
        return "invalid lobe at bounce %d -> non-unit direction -> replay weight eval/pdf of a later vertex not finite" % invalid[0]
    if invalid:
        return "invalid lobe at bounce %d, first non-finite factor %s" % (invalid[0], rec["first_nonfinite_factor"])
    if nonunit_in:
        return "non-unit wi without a flagged lobe, first non-finite factor %s" % rec["first_nonfinite_factor"]
    return "other: first non-finite factor %s" % rec["first_nonfinite_factor"]


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("scene", help="builtin scene name (scenes.load) or a path to a scene XML")
    ap.add_argument("--res", default="128")
    ap.add_argument("--spp", type=int, default=16)
    ap.add_argument("--seed", type=int, default=0)
    ap.add_argument("--sky", action="store_true", help="replace a missing envmap file by the procedural sky (scenes/disk)")
    ap.add_argument("--records", type=int, default=5, help="raw records to print")
    a = ap.parse_args()
    rx, _, ry = a.res.partition("x")
    rx, ry = int(rx), int(ry or rx)
    trace = tempfile.NamedTemporaryFile(suffix=".jsonl", delete=False).name
    os.environ["PLT_ORACLE_NAN_TRACE"] = trace
    import numpy as np
    import mitsuba3_plt_b200 as mi
    import oracle_lib
    from mitsuba3_plt_b200 import scenes
    if os.path.exists(a.scene):
        path = a.scene
        if a.sky:   # scenes/disk lacks its envmap blob (textures/empty_play_room.exr): copy the scene and supply a procedural sky
            import shutil
            from mitsuba3_plt_b200 import imageio
            d = os.path.join(tempfile.mkdtemp(), "scene")
            shutil.copytree(os.path.dirname(os.path.abspath(path)), d)
            os.chmod(os.path.join(d, "textures"), 0o755)
            imageio.write_exr(os.path.join(d, "textures", "empty_play_room.exr"), scenes.procedural_sky(64, 32), compression="zip")
            path = os.path.join(d, os.path.basename(path))
        scene = mi.load_file(path, resx=rx, resy=ry)
    else:
        scene = scenes.load(a.scene, resx=rx, resy=ry)
    r = oracle_lib.render(scene.flat, a.spp, seed=a.seed, max_depth=7, rr_depth=50, dump=True, n_threads=1)
    L = r["radiance"]
    bad = ~np.isfinite(L).all(axis=1)
    img_bad = ~np.isfinite(r["image"]).all(axis=-1)
    import re   # printf writes nan / inf; JSON spells them NaN / Infinity
    fix = lambda l: re.sub(r"(?<![\w\"])-?nan", "NaN", re.sub(r"(?<![\w\"])(-?)inf", r"\1Infinity", l))
    recs = [json.loads(fix(l)) for l in open(trace)]
    os.unlink(trace)
    print(f"scene {a.scene}  {rx}x{ry} x {a.spp} spp, seed {a.seed}, plt max_depth=7 rr_depth=50 (oracle, unpolarized spectral)")
    print(f"samples {len(L)}  non-finite samples {int(bad.sum())} ({bad.mean() * 1e6:.1f} per 1e6)  "
          f"non-finite pixels {int(img_bad.sum())} of {img_bad.size} ({img_bad.mean() * 100:.2f} %)")
    print(f"non-finite terms traced: {len(recs)}")
    by = collections.Counter(classify(x) for x in recs)
    for k, v in by.most_common():
        print(f"  {v:6d}  {k}")
    fac = collections.Counter((x["term"], x["first_nonfinite_factor"]) for x in recs)
    print("by (term, first non-finite factor):", dict((f"{t}/{f}", n) for (t, f), n in fac.items()))
    wn = [x["path"][-1]["wi_norm"] for x in recs]
    if wn:
        print(f"|wi| at the failing vertex: min {min(wn):.3f} median {sorted(wn)[len(wn) // 2]:.3f} max {max(wn):.3f}")
    for x in recs[:a.records]:
        print(json.dumps(x))


if __name__ == "__main__":
    main()
