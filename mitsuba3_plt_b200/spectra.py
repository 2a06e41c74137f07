"""Host-side (load-time) spectral helpers: CIE tables, rgb2spec coefficient lookup, sRGB decoding.

Mirrors, for the spectral variants:
  * srgb_model_fetch -> rgb2spec_fetch   (src/render/srgb.cpp:14-36, ext/rgb2spec/rgb2spec.c:70-130)
  * CIE1932Tables::initialize            (include/mitsuba/core/spectrum.h:160-182)
  * srgb_model_mean                      (include/mitsuba/render/srgb.h:25-36)
`data/srgb.coeff` is the reference's build product (`rgb2spec_opt 64 srgb.coeff`, ext/rgb2spec/CMakeLists.txt:28-32),
regenerated with tools/gen_srgb_coeff.sh and shipped as a data file.
"""
import os
import struct
import numpy as np

_DATA = os.path.join(os.path.dirname(__file__), "data")

MI_CIE_MIN, MI_CIE_MAX, MI_CIE_SAMPLES = 360.0, 830.0, 95
MI_CIE_D65_NORMALIZATION = 1.0 / 98.99741751876255
MI_CIE_Y_NORMALIZATION = 1.0 / 106.7502593994140625

_tables = None
_model = None


def cie_tables():
    """returns dict(xyz=(3,95) f32, srgb=(3,95) f32, d65=(95,) f32)"""
    global _tables
    if _tables is None:
        z = np.load(os.path.join(_DATA, "cie_tables.npz"))
        xyz = z["xyz"].astype(np.float32)
        # xyz_to_srgb (include/mitsuba/core/spectrum.h:406-412), float32 fmadd chain over columns
        M = np.array([[3.240479, -1.537150, -0.498535],
                      [-0.969256, 1.875991, 0.041556],
                      [0.055648, -0.204043, 1.057311]], dtype=np.float32)
        srgb = np.empty_like(xyz)
        for r in range(3):
            acc = (M[r, 0] * xyz[0]).astype(np.float32)
            for c in (1, 2):
                acc = (np.float64(M[r, c]) * xyz[c].astype(np.float64) + acc.astype(np.float64)).astype(np.float32)
            srgb[r] = acc
        _tables = dict(xyz=xyz, srgb=np.ascontiguousarray(srgb), d65=z["d65"].astype(np.float32))
    return _tables


def _load_model():
    global _model
    if _model is None:
        path = os.path.join(_DATA, "srgb.coeff")
        with open(path, "rb") as f:
            if f.read(4) != b"SPEC":
                raise RuntimeError("Could not load sRGB-to-spectrum upsampling model ('data/srgb.coeff')")
            res, = struct.unpack("<I", f.read(4))
            scale = np.frombuffer(f.read(4 * res), dtype=np.float32)
            data = np.frombuffer(f.read(4 * res ** 3 * 9), dtype=np.float32)
        _model = (res, scale.copy(), data.copy())
    return _model


def srgb_model_fetch(rgb):
    """rgb: (..., 3) array-like in [0,1] -> (..., 3) float32 coefficients (vectorised rgb2spec_fetch)."""
    res, scale, data = _load_model()
    rgb = np.clip(np.asarray(rgb, dtype=np.float32), 0.0, 1.0)
    shape = rgb.shape
    rgb = rgb.reshape(-1, 3)
    out = np.zeros_like(rgb)
    mono = (rgb[:, 0] == rgb[:, 1]) & (rgb[:, 1] == rgb[:, 2])
    # monochromatic: analytic
    v = rgb[mono, 0]
    with np.errstate(divide="ignore", invalid="ignore"):
        r = ((v - np.float32(.5)) / np.sqrt(v * (np.float32(1) - v))).astype(np.float32)
    r = np.where(v == 0, -np.inf, np.where(v == 1, np.inf, r)).astype(np.float32)
    out[mono, 2] = r
    idx = np.nonzero(~mono)[0]
    if idx.size:
        c = rgb[idx]
        i = np.zeros(len(c), dtype=np.int64)
        for j in (1, 2):
            i = np.where(c[np.arange(len(c)), j] >= c[np.arange(len(c)), i], j, i)
        ar = np.arange(len(c))
        z = c[ar, i]
        sc = (np.float32(res - 1) / z).astype(np.float32)
        x = (c[ar, (i + 1) % 3] * sc).astype(np.float32)
        y = (c[ar, (i + 2) % 3] * sc).astype(np.float32)
        xi = np.minimum(x.astype(np.uint32), res - 2).astype(np.int64)
        yi = np.minimum(y.astype(np.uint32), res - 2).astype(np.int64)
        # rgb2spec_find_interval: largest zi with scale[zi] <= z, clamped to res-2
        zi = np.clip(np.searchsorted(scale, z, side="right") - 1, 0, res - 2).astype(np.int64)
        offset = (((i * res + zi) * res + yi) * res + xi) * 3
        dx, dy, dz = 3, 3 * res, 3 * res * res
        x1 = (x - xi.astype(np.float32)).astype(np.float32); x0 = (np.float32(1) - x1)
        y1 = (y - yi.astype(np.float32)).astype(np.float32); y0 = (np.float32(1) - y1)
        z1 = ((z - scale[zi]) / (scale[zi + 1] - scale[zi])).astype(np.float32); z0 = (np.float32(1) - z1)
        f32 = np.float32
        for j in range(3):
            o = offset + j
            a = ((data[o] * x0 + data[o + dx] * x1).astype(f32) * y0 +
                 (data[o + dy] * x0 + data[o + dy + dx] * x1).astype(f32) * y1).astype(f32) * z0
            b = ((data[o + dz] * x0 + data[o + dz + dx] * x1).astype(f32) * y0 +
                 (data[o + dz + dy] * x0 + data[o + dz + dy + dx] * x1).astype(f32) * y1).astype(f32) * z1
            out[idx, j] = (a + b).astype(f32)
    return out.reshape(shape)


def srgb_model_eval(coeff, wavelengths):
    """float32 evaluation (include/mitsuba/render/srgb.h:9-23); used for means and host-side checks."""
    c = np.asarray(coeff, dtype=np.float32)
    w = np.asarray(wavelengths, dtype=np.float32)
    v = (c[..., 0:1] * w + c[..., 1:2]) * w + c[..., 2:3]
    with np.errstate(invalid="ignore"):
        r = np.maximum(0, .5 * v / np.sqrt(v * v + 1) + .5)
    inf = np.isinf(c[..., 2:3])
    return np.where(inf, np.sign(c[..., 2:3]) * .5 + .5, r).astype(np.float32)


def srgb_model_mean(coeff):
    lam = np.linspace(MI_CIE_MIN, MI_CIE_MAX, 16, dtype=np.float32)
    return srgb_model_eval(np.asarray(coeff, dtype=np.float32)[..., None, :], lam).mean(axis=-1).astype(np.float32)


def srgb_to_linear(x):
    """Bitmap::convert sRGB gamma removal for 8-bit inputs (src/core/bitmap.cpp, struct converter)."""
    x = np.asarray(x, dtype=np.float32)
    return np.where(x <= 0.04045, x / 12.92, ((x + 0.055) / 1.055) ** 2.4).astype(np.float32)


def xyz_to_srgb(xyz):
    """include/mitsuba/core/spectrum.h:406-412 (float32)"""
    M = np.array([[3.240479, -1.537150, -0.498535], [-0.969256, 1.875991, 0.041556], [0.055648, -0.204043, 1.057311]], dtype=np.float32)
    x = np.asarray(xyz, dtype=np.float32)
    acc = M[:, 0] * x[0]
    for c in (1, 2):
        acc = (M[:, c].astype(np.float64) * np.float64(x[c]) + acc.astype(np.float64)).astype(np.float32)
    return acc.astype(np.float32)


def spectrum_list_to_srgb(wavelengths, values, bounded=True, d65=False):
    """src/core/spectrum.cpp:85-130 (float32): integrate a piecewise-linear spectrum against the CIE curves (times D65 for
    reflectances) in 1000 steps and convert to linear sRGB; what the RGB variants do with <spectrum value="w:v, ..."> at load time."""
    F = np.float32
    wl = np.asarray(wavelengths, dtype=F); vals = np.asarray(values, dtype=F)
    tabs = cie_tables()
    steps = 1000
    xyz = np.zeros(3, dtype=F)
    scale = F(MI_CIE_SAMPLES - 1) / (F(MI_CIE_MAX) - F(MI_CIE_MIN))
    for i in range(steps):
        w = F(MI_CIE_MIN) + (F(i) / F(steps - 1)) * (F(MI_CIE_MAX) - F(MI_CIE_MIN))
        if w < wl[0] or w > wl[-1]:
            continue
        index = int(np.clip(np.searchsorted(wl, w, side="right") - 1, 0, len(wl) - 2))   # math::find_interval
        w0, w1, v0, v1 = wl[index], wl[index + 1], vals[index], vals[index + 1]
        v = ((w - w1) * v0 + (w0 - w) * v1) / (w0 - w1)
        t = (w - F(MI_CIE_MIN)) * scale
        i0 = int(min(max(int(t), 0), MI_CIE_SAMPLES - 2))
        f1 = F(t - F(i0)); f0 = F(1) - f1
        c = (f0 * tabs["xyz"][:, i0] + f1 * tabs["xyz"][:, i0 + 1]).astype(F)
        dv = F(1)
        if d65:   # cie_d65(w): table lookup times MI_CIE_D65_NORMALIZATION (spectrum.h:300-326)
            dv = F((f0 * tabs["d65"][i0] + f1 * tabs["d65"][i0 + 1]) * F(MI_CIE_D65_NORMALIZATION))
        xyz = (xyz + c * v * dv).astype(F)
    xyz = (xyz * ((F(MI_CIE_MAX) - F(MI_CIE_MIN)) / F(steps))).astype(F)
    rgb = xyz_to_srgb(xyz)
    return np.clip(rgb, 0, 1).astype(F) if bounded else np.maximum(rgb, 0).astype(F)
