"""The C-ABI library loads and exports every symbol include/pltb.h declares (no compute calls: runs without a GPU)."""
import ctypes as C
import os
import re

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_exports_match_header():
    from mitsuba3_plt_b200 import capi
    hdr = open(os.path.join(ROOT, "include", "pltb.h")).read()
    declared = set(re.findall(r"\b(pltb_[a-z0-9_]+)\s*\(", hdr))
    declared -= {"pltb_bessel_tables"}
    assert declared == set(capi.EXPORTS), declared ^ set(capi.EXPORTS)
    L = capi.lib()
    for name in declared:
        assert getattr(L, name) is not None
    assert L.pltb_abi_version() == capi.PLTB_ABI_VERSION


def test_struct_sizes_match_header():
    """ctypes mirrors must have the C layout: compile a tiny probe with gcc against the header"""
    import subprocess, tempfile
    from mitsuba3_plt_b200 import capi
    names = {"pltb_texture": capi.Texture, "pltb_bsdf": capi.Bsdf, "pltb_shape": capi.Shape, "pltb_emitter": capi.Emitter,
             "pltb_sensor": capi.Sensor, "pltb_film": capi.Film, "pltb_scene_desc": capi.SceneDesc,
             "pltb_render_params": capi.RenderParams, "pltb_sample_dump": capi.SampleDump, "pltb_stats": capi.Stats}
    src = '#include <stdio.h>\n#include "pltb.h"\nint main(){' + "".join(
        f'printf("{n} %zu\\n", sizeof({n}));' for n in names) + "return 0;}"
    with tempfile.TemporaryDirectory() as d:
        open(os.path.join(d, "p.c"), "w").write(src)
        subprocess.check_call(["gcc", "-I", os.path.join(ROOT, "include"), "-o", os.path.join(d, "p"), os.path.join(d, "p.c")])
        out = subprocess.check_output([os.path.join(d, "p")]).decode().split()
    sizes = dict(zip(out[::2], map(int, out[1::2])))
    for n, cls in names.items():
        assert C.sizeof(cls) == sizes[n], (n, C.sizeof(cls), sizes[n])


def test_no_device_fails_loudly():
    """no CPU fallback: without a CUDA device context creation returns an error, it does not render on the host"""
    import torch
    if torch.cuda.is_available():
        return
    from mitsuba3_plt_b200 import capi
    h = C.c_void_p()
    rc = capi.lib().pltb_context_create(-1, 0, 1, None, C.byref(h))
    assert rc == -2 and b"no CPU fallback" in capi.lib().pltb_last_error()
