#!/usr/bin/env python
"""Per-sample agreement of the CUDA path with the oracle on the cases where the 99.9 % bar is tightest, for the library selected
by PLTB200_LIB (build variants: tools/build_variant.sh).  One JSON line per case; used to price numerics choices (MUFU.RSQ,
reduced-precision Bessel in the lobe selection) in divergent samples.
usage: python tools/parity_ab.py [case ...]   cases: crop c1 c2 veach"""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT); sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np
import mitsuba3_plt_b200 as mi
from mitsuba3_plt_b200 import scenes
import oracle_lib
from parity import nonfinite_counts, sample_agreement

CASES = {"crop": ("cbox-grating", 1024, 1024, 0, (316, 656, 16, 16)), "c1": ("cbox", 256, 16, 0, None),
         "c2": ("cbox-grating", 256, 16, 1, None), "veach": ("veach-mis", 256, 16, 1, None)}
it = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
for c in (sys.argv[1:] or list(CASES)):
    name, res, spp, seed, crop = CASES[c]
    sc = scenes.load(name, resx=res, resy=res)
    dump = {}
    it.render(sc, spp=spp, seed=seed, dump=dump, crop=crop)
    ref = oracle_lib.render(sc.flat, spp, seed=seed, dump=True, crop=crop, film=False)
    a, b = dump["radiance"].astype(np.float64), ref["radiance"].astype(np.float64)
    with np.errstate(invalid="ignore", divide="ignore"):
        err = (np.abs(a - b) / np.maximum(np.abs(b), 1e-6)).max(axis=1)
    both_nan = (np.isnan(a) & np.isnan(b)).any(axis=1)
    err[both_nan] = 0
    bad = ~(err <= 1e-4)
    rays_differ = (dump["n_rays"] != ref["n_rays"]).any(axis=1)
    print(json.dumps({"lib": os.environ.get("PLTB200_LIB", "default"), "case": c, "samples": len(a),
                      "agreement": sample_agreement(a, b), "divergent": int(bad.sum()),
                      "divergent_with_other_ray_counts": int((bad & rays_differ).sum()),
                      "divergent_err_1e-4..1e-3": int((bad & (err <= 1e-3)).sum()),
                      "divergent_err_1e-3..1e-1": int((bad & (err > 1e-3) & (err <= 1e-1)).sum()),
                      "divergent_err_gt_1e-1_or_nan": int((bad & ~(err <= 1e-1)).sum()),
                      "nonfinite_gpu_oracle_both": nonfinite_counts(a, b)}))
