"""Plugin instantiation and flattening: Properties tree -> host SoA buffers -> `pltb_scene_desc`.

Host-side mirror of what the reference does at scene-load time: the plugin constructors
(src/shapes/ply.cpp, src/bsdfs/*.cpp, src/emitters/area.cpp, src/sensors/perspective.cpp, src/films/hdrfilm.cpp,
src/spectra/*.cpp, src/textures/bitmap.cpp) and Scene's ctor (src/render/scene.cpp:22-96).  Nothing here runs per
sample.
"""
import ctypes as C
import math
import os

import numpy as np

from . import capi, spectra
from .xmlscene import Color, Properties, SpectrumValue

F32 = np.float32


def _m16(m):
    return (C.c_float * 16)(*np.asarray(m, dtype=np.float64).reshape(16).astype(F32))


# ---------------------------------------------------------------------------------------------------------------------
# PLY loader (src/shapes/ply.cpp): ascii / binary little-endian, vertex x,y,z[,nx,ny,nz][,s|u,t|v], face lists
# ---------------------------------------------------------------------------------------------------------------------
_PLY_T = {"char": "i1", "uchar": "u1", "short": "i2", "ushort": "u2", "int": "i4", "uint": "u4", "float": "f4",
          "double": "f8", "int8": "i1", "uint8": "u1", "int16": "i2", "uint16": "u2", "int32": "i4", "uint32": "u4",
          "float32": "f4", "float64": "f8"}


def load_ply(path):
    with open(path, "rb") as f:
        if f.readline().strip() != b"ply":
            raise RuntimeError(f'"{path}": invalid PLY header')
        fmt, elements = None, []
        while True:
            line = f.readline().decode().strip()
            if line == "end_header":
                break
            tok = line.split()
            if not tok or tok[0] in ("comment", "obj_info"):
                continue
            if tok[0] == "format":
                fmt = tok[1]
            elif tok[0] == "element":
                elements.append((tok[1], int(tok[2]), []))
            elif tok[0] == "property":
                elements[-1][2].append(tok[1:])
        data = {}
        for name, count, props in elements:
            if name == "face" or any(p[0] == "list" for p in props):
                lp = [p for p in props if p[0] == "list"][0]
                ct, it = _PLY_T[lp[1]], _PLY_T[lp[2]]
                if fmt == "ascii":
                    rows = [[int(x) for x in f.readline().split()] for _ in range(count)]
                    if any(r[0] != 3 for r in rows):
                        raise RuntimeError("ply: only triangle faces are supported")
                    data[name] = np.array([r[1:4] for r in rows], dtype=np.uint32)
                else:
                    end = "<" if "little" in fmt else ">"
                    rec = np.dtype([("n", end + ct), ("v", end + it, (3,))])
                    arr = np.frombuffer(f.read(rec.itemsize * count), dtype=rec)
                    if np.any(arr["n"] != 3):
                        raise RuntimeError("ply: only triangle faces are supported")
                    data[name] = arr["v"].astype(np.uint32)
            else:
                if fmt == "ascii":
                    arr = np.array([[float(x) for x in f.readline().split()] for _ in range(count)])
                    data[name] = {p[1]: arr[:, i] for i, p in enumerate(props)}
                else:
                    end = "<" if "little" in fmt else ">"
                    rec = np.dtype([(p[1], end + _PLY_T[p[0]]) for p in props])
                    arr = np.frombuffer(f.read(rec.itemsize * count), dtype=rec)
                    data[name] = {p[1]: arr[p[1]] for p in props}
    v = data["vertex"]
    pos = np.stack([v["x"], v["y"], v["z"]], 1).astype(F32)
    nrm = np.stack([v["nx"], v["ny"], v["nz"]], 1).astype(F32) if "nx" in v else None
    uv = None
    for a, b in (("s", "t"), ("u", "v"), ("texture_u", "texture_v")):
        if a in v:
            uv = np.stack([v[a], v[b]], 1).astype(F32)
            break
    return pos, nrm, uv, data["face"]


# ---------------------------------------------------------------------------------------------------------------------
# .serialized loader (src/shapes/serialized.cpp:240-400): header 0x041C + version 3/4, one zlib stream per sub-mesh, offset
# dictionary (uint64 for v4, uint32 for v3) followed by the mesh count at the end of the file
# ---------------------------------------------------------------------------------------------------------------------
_SER_HEADER, _SER_NORMALS, _SER_UV, _SER_COLORS, _SER_DOUBLE = 0x041C, 0x0001, 0x0002, 0x0008, 0x2000


def load_serialized(path, shape_index=0):
    import struct
    import zlib
    buf = open(path, "rb").read()
    fmt, version = struct.unpack_from("<HH", buf, 0)
    if fmt != _SER_HEADER:
        raise RuntimeError(f'"{path}": encountered an invalid file format!')
    if version not in (3, 4):
        raise RuntimeError(f'"{path}": encountered an incompatible file version!')
    start = 4
    if shape_index != 0:
        count, = struct.unpack_from("<I", buf, len(buf) - 4)
        if shape_index > count:
            raise RuntimeError(f"Unable to unserialize mesh, shape index is out of range! (requested {shape_index} out of "
                               f"0..{count - 1})")
        if version == 4:
            offset, = struct.unpack_from("<Q", buf, len(buf) - 8 * (count - shape_index) - 4)
        else:
            offset, = struct.unpack_from("<I", buf, len(buf) - 4 * (count - shape_index + 1))
        start = offset + 4
    raw = zlib.decompressobj().decompress(buf[start:])
    flags, = struct.unpack_from("<I", raw, 0)
    at = 4
    if version == 4:
        at = raw.index(b"\0", at) + 1
    nv, nf = struct.unpack_from("<QQ", raw, at)
    at += 16
    ft = np.dtype("<f8" if flags & _SER_DOUBLE else "<f4")

    def take(dim):
        nonlocal at
        a = np.frombuffer(raw, ft, nv * dim, at).reshape(nv, dim).astype(F32)
        at += nv * dim * ft.itemsize
        return a
    pos = take(3)
    nrm = take(3) if flags & _SER_NORMALS else None
    uv = take(2) if flags & _SER_UV else None
    if flags & _SER_COLORS:
        take(3)
    faces = np.frombuffer(raw, "<u4", nf * 3, at).reshape(nf, 3).astype(np.uint32)
    return pos, nrm, uv, faces


def write_serialized(path, meshes, version=4):
    """The inverse of load_serialized, for fixtures: meshes = [(positions, normals | None, texcoords | None, faces), ...]"""
    import struct
    import zlib
    out, offsets = bytearray(), []
    for i, (pos, nrm, uv, faces) in enumerate(meshes):
        offsets.append(len(out))
        flags = 0x1000 | (_SER_NORMALS if nrm is not None else 0) | (_SER_UV if uv is not None else 0)
        body = struct.pack("<I", flags) + (b"mesh%d\0" % i if version == 4 else b"")
        body += struct.pack("<QQ", len(pos), len(faces)) + np.asarray(pos, "<f4").tobytes()
        if nrm is not None:
            body += np.asarray(nrm, "<f4").tobytes()
        if uv is not None:
            body += np.asarray(uv, "<f4").tobytes()
        body += np.asarray(faces, "<u4").tobytes()
        out += struct.pack("<HH", _SER_HEADER, version) + zlib.compress(body)
    for o in offsets:
        out += struct.pack("<Q" if version == 4 else "<I", o)
    out += struct.pack("<I", len(meshes))
    with open(path, "wb") as f:
        f.write(bytes(out))


def _xf_points(m, p):
    m = np.asarray(m, dtype=np.float64)
    return (p.astype(np.float64) @ m[:3, :3].T + m[:3, 3]).astype(F32)


def _xf_normals(m, n):
    it = np.linalg.inv(np.asarray(m, dtype=np.float64)[:3, :3]).T
    r = n.astype(np.float64) @ it.T
    r /= np.linalg.norm(r, axis=1, keepdims=True)
    return r.astype(F32)


def _cube_mesh():
    """src/shapes/cube.cpp: 24 vertices / 12 triangles, one quad per face with its own normal and [0,1]^2 uvs."""
    pos, nrm, uv, faces = [], [], [], []
    for axis in range(3):
        for sgn in (-1.0, 1.0):
            n = np.zeros(3); n[axis] = sgn
            a = np.zeros(3); a[(axis + 1) % 3] = 1.0
            b = np.cross(n, a)
            base = len(pos)
            for (s, t) in ((-1, -1), (1, -1), (1, 1), (-1, 1)):
                pos.append(n + s * a + t * b); nrm.append(n); uv.append(((s + 1) / 2, (t + 1) / 2))
            faces += [(base, base + 1, base + 2), (base + 3, base, base + 2)]
    return (np.array(pos, dtype=F32), np.array(nrm, dtype=F32), np.array(uv, dtype=F32),
            np.array(faces, dtype=np.uint32))


# ---------------------------------------------------------------------------------------------------------------------
# Flattener
# ---------------------------------------------------------------------------------------------------------------------
class FlatScene:
    """Owns the numpy buffers behind a `pltb_scene_desc` (keeps them alive) plus the traversable parameters."""

    def __init__(self, rgb=False):
        # rgb: flatten for the `*_rgb` variants (Spectrum = Color3f): every spectrum / texture becomes an RGB value
        # (PLTB_TEX_RGB, PLTB_TEX_UNIFORM or an RGB bitmap), as the plugins do under is_rgb_v<Spectrum>
        self.rgb = bool(rgb)
        self.textures, self.tex_data = [], []
        self.bsdfs, self.shapes, self.emitters, self.sensors = [], [], [], []
        self.positions, self.normals, self.texcoords, self.faces, self.face_shape = [], [], [], [], []
        self.area_cdf = []
        self.n_vertices = 0
        self.n_tex_floats = 0
        self.n_cdf = 0
        self.film = None
        self._bsdf_cache = {}
        self.spp = 4
        self.integrator_props = None
        self.sensor_ids = []
        self.bbox = [np.full(3, np.inf), np.full(3, -np.inf)]

    # ---- textures / spectra ------------------------------------------------------------------------------------------
    def _push_data(self, arr):
        arr = np.ascontiguousarray(arr, dtype=F32).reshape(-1)
        off = self.n_tex_floats
        self.tex_data.append(arr)
        self.n_tex_floats += arr.size
        return off

    def _add_tex(self, **kw):
        t = capi.Texture()
        t.to_uv[:] = [1, 0, 0, 0, 1, 0, 0, 0, 1]
        for k, v in kw.items():
            if k == "p":
                for i, x in enumerate(v):
                    t.p[i] = x
            elif k == "to_uv":
                t.to_uv[:] = list(np.asarray(v, dtype=F32).reshape(9))
            else:
                setattr(t, k, v)
        self.textures.append(t)
        return len(self.textures) - 1

    def d65_table(self, scale):
        tab = spectra.cie_tables()["d65"].astype(np.float64) * float(F32(scale)) * F32(spectra.MI_CIE_D65_NORMALIZATION)
        return tab.astype(F32)

    def texture(self, props, name, default, within_emitter=False):
        """props.texture<Texture>(name, default) incl. the rgb/spectrum -> plugin rules of xml.cpp:1146-1265.
        `default` None means "absent -> -1"."""
        v = props.get(name) if props is not None else None
        if v is None:
            if default is None:
                return -1
            v = float(default)
        if isinstance(v, (int, float)) and not isinstance(v, bool):
            return self._add_tex(type=capi.TEX_UNIFORM, p=[float(v)], mean=float(v))
        if self.rgb:
            return self._texture_rgb(v, name, within_emitter)
        if isinstance(v, Color):
            rgb = np.array(v, dtype=F32)
            if within_emitter:  # src/spectra/d65.cpp:74-112
                scale = F32(1.0)
                factor = F32(rgb.max() * F32(2))
                if factor != 0:
                    rgb = (rgb / factor).astype(F32)
                scale = F32(scale * factor)
                coeff = spectra.srgb_model_fetch(rgb)
                off = self._push_data(self.d65_table(scale))
                return self._add_tex(type=capi.TEX_D65, flags=1, offset=off, count=95,
                                     p=[coeff[0], coeff[1], coeff[2], 360.0, 830.0])
            # xml.cpp:49-51,1153-1154: eta / k / int_ior / ext_ior become `srgb` with unbounded=true (no range check; the
            # coefficient fetch itself still clamps to [0, 1], ext/rgb2spec/rgb2spec.c:80-82)
            unbounded = name in ("eta", "k", "int_ior", "ext_ior")
            if (np.any(rgb < 0) or np.any(rgb > 1)) and not unbounded:
                raise RuntimeError(f"Invalid RGB reflectance value {list(rgb)}, must be in the range [0, 1]!")
            coeff = spectra.srgb_model_fetch(rgb)
            return self._add_tex(type=capi.TEX_SRGB, p=list(coeff), mean=float(np.asarray(spectra.srgb_model_mean(coeff)).reshape(-1)[0]))
        if isinstance(v, SpectrumValue):
            if v.const is not None:
                if within_emitter:
                    pass  # spectral mode: plain uniform (xml.cpp:1176-1190)
                return self._add_tex(type=capi.TEX_UNIFORM, p=[v.const], mean=v.const)
            return self.spectrum_table(v.wavelengths, v.values)
        if isinstance(v, Properties):
            return self.texture_plugin(v, within_emitter)
        raise RuntimeError(f'property "{name}": unsupported value {v!r}')

    # ---- RGB variants (is_rgb_v<Spectrum>) ---------------------------------------------------------------------------
    def _rgb_tex(self, rgb):
        rgb = np.asarray(rgb, dtype=F32)
        lum = float(F32(rgb[0]) * F32(0.212671) + F32(rgb[1]) * F32(0.715160) + F32(rgb[2]) * F32(0.072169))
        return self._add_tex(type=capi.TEX_RGB, p=[float(rgb[0]), float(rgb[1]), float(rgb[2])], mean=lum)

    def _texture_rgb(self, v, name, within_emitter):
        """The rgb / spectrum -> plugin rules of xml.cpp:1146-1260 and the plugin constructors in an RGB variant."""
        unbounded = name in ("eta", "k", "int_ior", "ext_ior")
        if isinstance(v, Color):
            # within an emitter: d65 with `color`, which expands to srgb(color * scale, unbounded) (d65.cpp:100-156); else srgb
            rgb = np.array(v, dtype=F32)
            if not within_emitter and not unbounded and (np.any(rgb < 0) or np.any(rgb > 1)):
                raise RuntimeError(f"Invalid RGB reflectance value {list(rgb)}, must be in the range [0, 1]!")
            return self._rgb_tex(rgb)
        if isinstance(v, SpectrumValue):
            if v.const is not None:
                if within_emitter and not unbounded:   # xml.cpp:1176-1188: const * xyz_to_srgb(Color3f(1))
                    white = spectra.xyz_to_srgb(np.ones(3, dtype=F32))
                    return self._rgb_tex((F32(v.const) * white).astype(F32))
                return self._add_tex(type=capi.TEX_UNIFORM, p=[v.const], mean=v.const)
            vals = np.asarray(v.values, dtype=F32) * F32(spectra.MI_CIE_Y_NORMALIZATION)   # xml.cpp:1200-1204
            rgb = spectra.spectrum_list_to_srgb(np.asarray(v.wavelengths, dtype=F32), vals,
                                                bounded=not (within_emitter or unbounded), d65=not (within_emitter and not unbounded))
            return self._rgb_tex(rgb)
        if isinstance(v, Properties):
            return self.texture_plugin(v, within_emitter)
        raise RuntimeError(f'property "{name}": unsupported value {v!r}')

    def spectrum_table(self, wl, vals):
        """regular / irregular spectrum (xml.cpp:1192-1232); the CDF used by sample_spectrum is appended."""
        wl = np.asarray(wl, dtype=np.float64); vals = np.asarray(vals, dtype=np.float64)
        iv = np.diff(wl)
        regular = (iv.max() - iv.min()) < 1e-3 * iv.min()
        wl32, v32 = wl.astype(F32), vals.astype(F32)
        n = len(wl)
        if regular:
            isz = F32((wl32[-1] - wl32[0]) / F32(n - 1))
            # ContinuousDistribution::compute_cdf (distr_1d.h:540-585): trapezoid sums in double
            cdf = np.cumsum(0.5 * float(isz) * (v32[1:].astype(np.float64) + v32[:-1])).astype(F32)
            off = self._push_data(np.concatenate([v32, cdf, [0]]))
            return self._add_tex(type=capi.TEX_REGULAR, offset=off, count=n, p=[0, 0, 0, wl32[0], wl32[-1]])
        cdf = np.cumsum(0.5 * np.diff(wl32.astype(np.float64)) * (v32[1:].astype(np.float64) + v32[:-1])).astype(F32)
        off = self._push_data(np.concatenate([wl32, v32, cdf, [0]]))
        return self._add_tex(type=capi.TEX_IRREGULAR, offset=off, count=n, p=[0, 0, 0, wl32[0], wl32[-1]])

    def texture_plugin(self, p, within_emitter=False):
        t = p.plugin_name
        if t == "uniform":
            v = float(p.get("value", 1.0))
            return self._add_tex(type=capi.TEX_UNIFORM, p=[v], mean=v)
        if self.rgb:
            if t == "srgb":   # srgb.cpp:52-57
                rgb = np.array(p["color"], dtype=F32)
                if (np.any(rgb < 0) or np.any(rgb > 1)) and not p.get("unbounded", False):
                    raise RuntimeError(f"Invalid RGB reflectance value {list(rgb)}, must be in the range [0, 1]!")
                return self._rgb_tex(rgb)
            if t == "d65":    # expands to srgb(color * scale, unbounded) or uniform(scale) (d65.cpp:140-160)
                scale = F32(p.get("scale", 1.0))
                if "color" in p:
                    return self._rgb_tex((np.array(p["color"], dtype=F32) * scale).astype(F32))
                return self._add_tex(type=capi.TEX_UNIFORM, p=[float(scale)], mean=float(scale))
            if t == "bitmap":
                return self.bitmap(p)
            raise RuntimeError(f'Plugin "{t}": spectral textures are not available in the RGB variants (use <rgb> / <spectrum value=...>)')
        if t == "srgb":
            return self.texture(Properties("", color=Color(p["color"])), "color", None)
        if t == "d65":
            scale = float(p.get("scale", 1.0))
            if "color" in p:
                q = Properties("", c=Color(np.asarray(p["color"]) * scale))
                return self.texture(q, "c", None, within_emitter=True)
            off = self._push_data(self.d65_table(scale))
            return self._add_tex(type=capi.TEX_D65, flags=0, offset=off, count=95, p=[0, 0, 0, 360.0, 830.0])
        def numbers(v):   # "1, 2" / "1 2" strings as well as sequences (src/spectra/regular.cpp:38-47 tokenises on " ,")
            return np.array([float(x) for x in v.replace(",", " ").split()] if isinstance(v, str) else v, dtype=np.float64)
        if t == "regular":
            vals = numbers(p["values"])
            wl = np.linspace(float(p["wavelength_min"]), float(p["wavelength_max"]), len(vals))
            return self.spectrum_table(wl, vals)
        if t == "irregular":
            wl, vals = numbers(p["wavelengths"]), numbers(p["values"])
            if len(wl) != len(vals):
                raise RuntimeError("IrregularSpectrum: 'wavelengths' and 'values' parameters must have the same size!")
            return self.spectrum_table(wl, vals)
        if t == "blackbody":
            return self._add_tex(type=capi.TEX_BLACKBODY,
                                 p=[float(p.get("temperature", 5000.0)), 0, 0,
                                    float(p.get("wavelength_min", 360.0)), float(p.get("wavelength_max", 830.0))])
        if t == "bitmap":
            return self.bitmap(p)
        raise RuntimeError(f'Plugin "{t}" not found (texture)')

    def bitmap(self, p):
        """src/textures/bitmap.cpp: 8-bit sRGB -> linear RGB float32 -> rgb2spec coefficients per texel (:321-333)"""
        from PIL import Image
        path = p["filename"]
        if not os.path.isabs(path):
            path = os.path.join(getattr(p, "base_path", "."), path)
        im = Image.open(path)
        raw = bool(p.get("raw", False))
        if im.mode in ("L", "I", "F", "1") and raw:
            raise RuntimeError("bitmap: single-channel textures are not on the plt path")
        a = np.asarray(im.convert("RGB"), dtype=F32) / F32(255)
        if not raw:
            a = spectra.srgb_to_linear(a)
        if self.rgb:   # RGB variants: the texels stay linear RGB (bitmap.cpp:230-250), mean = mean luminance
            coeff = np.ascontiguousarray(a, dtype=F32)
        else:
            coeff = spectra.srgb_model_fetch(a)
        h, w = coeff.shape[:2]
        wrap = {"repeat": capi.WRAP_REPEAT, "mirror": capi.WRAP_MIRROR, "clamp": capi.WRAP_CLAMP}[p.get("wrap_mode", "repeat")]
        if p.get("filter_type", "bilinear") != "bilinear":
            raise RuntimeError("bitmap: only bilinear filtering is supported")
        to_uv = np.eye(3)
        if "to_uv" in p:
            m = np.asarray(p["to_uv"], dtype=np.float64)
            to_uv = m if m.shape == (3, 3) else np.array([[m[0, 0], m[0, 1], m[0, 3]], [m[1, 0], m[1, 1], m[1, 3]], [0, 0, 1]])
        off = self._push_data(coeff)
        if self.rgb:
            c = coeff.reshape(-1, 3)
            mean = float((c[:, 0] * F32(0.212671) + c[:, 1] * F32(0.715160) + c[:, 2] * F32(0.072169)).mean())
        else:
            mean = float(spectra.srgb_model_mean(coeff.reshape(-1, 3)).mean())
        return self._add_tex(type=capi.TEX_BITMAP, flags=capi.TEXFLAG_RGB if self.rgb else 0, offset=off, count=w * h * 3,
                             width=w, height=h, wrap=wrap, to_uv=to_uv, mean=mean)

    def eval_1(self, tex_id):
        """Texture::eval_1 for spatially uniform parameters (roughness, grating parameters)."""
        t = self.textures[tex_id]
        if t.type == capi.TEX_UNIFORM:
            return float(t.p[0])
        if t.type in (capi.TEX_SRGB, capi.TEX_RGB):
            return float(t.mean)
        raise RuntimeError("spatially varying scalar BSDF parameters are not supported on the plt path")

    def scalar(self, props, name, default):
        v = props.get(name)
        if v is None:
            return float(default)
        if isinstance(v, (int, float)):
            return float(v)
        return self.eval_1(self.texture(props, name, default))

    # ---- BSDFs ---------------------------------------------------------------------------------------------------------
    def bsdf(self, p):
        key = id(p)
        if key in self._bsdf_cache:
            return self._bsdf_cache[key]
        b = capi.Bsdf()
        b.tex[:] = [-1, -1, -1, -1]
        flags = 0
        inner = p
        if p.plugin_name == "twosided":
            nested = [c for _, c in p.all_children()]
            if len(nested) != 1:
                raise RuntimeError("twosided: exactly one nested BSDF is supported on the plt path")
            inner, flags = nested[0], capi.BSDF_FLAG_TWOSIDED
            if inner.plugin_name in ("dielectric", "twosided"):
                raise RuntimeError("twosided: only materials without a transmission component can be nested!")
        t = inner.plugin_name
        b.flags = flags

        def conductor_ior():
            material = inner.get("material", "none")
            if material != "none" and "eta" not in inner:
                raise RuntimeError(f'conductor material "{material}": measured IOR files are not shipped; give eta/k')
            b.tex[0] = self.texture(inner, "eta", 0.0)
            b.tex[1] = self.texture(inner, "k", 1.0)

        def microfacet():
            d = str(inner.get("distribution", "beckmann")).lower()
            if d not in ("beckmann", "ggx"):
                raise RuntimeError(f'Specified an invalid distribution "{d}", must be "beckmann" or "ggx"!')
            b.u[0] = capi.MICROFACET_BECKMANN if d == "beckmann" else capi.MICROFACET_GGX
            b.u[1] = 1 if inner.get("sample_visible", True) else 0
            if "alpha_u" in inner or "alpha_v" in inner:
                if not ("alpha_u" in inner and "alpha_v" in inner):
                    raise RuntimeError("Microfacet model: both 'alpha_u' and 'alpha_v' must be specified.")
                if "alpha" in inner:
                    raise RuntimeError("Microfacet model: please specify either 'alpha' or 'alpha_u'/'alpha_v'.")
                b.f[1], b.f[2] = self.scalar(inner, "alpha_u", 0.1), self.scalar(inner, "alpha_v", 0.1)
            else:
                b.f[1] = b.f[2] = self.scalar(inner, "alpha", 0.1)

        if t == "diffuse":
            b.type = capi.BSDF_DIFFUSE
            b.tex[0] = self.texture(inner, "reflectance", 0.5)
        elif t == "conductor":
            b.type = capi.BSDF_CONDUCTOR
            conductor_ior()
            b.tex[2] = self.texture(inner, "specular_reflectance", 1.0)
        elif t == "dielectric":
            b.type = capi.BSDF_DIELECTRIC
            int_ior, ext_ior = _lookup_ior(inner.get("int_ior", "bk7")), _lookup_ior(inner.get("ext_ior", "air"))
            if int_ior < 0 or ext_ior < 0:
                raise RuntimeError("The interior and exterior indices of refraction must be positive!")
            b.f[0] = F32(int_ior) / F32(ext_ior)
            b.tex[2] = self.texture(inner, "specular_reflectance", None)
            b.tex[3] = self.texture(inner, "specular_transmittance", None)
        elif t == "roughconductor":
            b.type = capi.BSDF_ROUGHCONDUCTOR
            conductor_ior(); microfacet()
            b.tex[2] = self.texture(inner, "specular_reflectance", None)
        elif t == "roughgrating":  # src/bsdfs/roughgrating.cpp:180-276
            b.type = capi.BSDF_ROUGHGRATING
            conductor_ior(); microfacet()
            b.tex[2] = self.texture(inner, "specular_reflectance", None)
            b.f[3] = self.scalar(inner, "grating_angle", 0.0)
            if "inv_period_x" in inner:
                if "inv_period" in inner:
                    raise RuntimeError("Grating inv inv_period: please specify either 'inv_period' or "
                                       "'inv_period_x'/'inv_period_x'.")
                b.f[4], b.f[5] = self.scalar(inner, "inv_period_x", 0.1), self.scalar(inner, "inv_period_y", 0.1)
            else:
                b.f[4] = b.f[5] = self.scalar(inner, "inv_period", 0.1)
            b.f[6] = self.scalar(inner, "height", 0.3)
            b.f[7] = self.scalar(inner, "multiplier", 1.0)
            b.f[8] = float(inner.get("coherence", 1e-18))
            lobes = int(inner.get("lobes", 5))
            if lobes < 1 or lobes > 9:
                raise RuntimeError("roughgrating: 'lobes' must be in [1, 9]")
            b.u[2] = lobes
            lt = inner.get("lobe_type", "rectangular")
            types = {"sinusoidal": 0, "rectangular": 1, "linear": 2}
            if lt not in types:
                raise RuntimeError(f"Grating surface type {lt} not supported!")
            b.u[3] = types[lt] | (0x10 if inner.get("radial", False) else 0)
        elif t == "principled":  # src/bsdfs/principled.cpp:180-236
            b.type = capi.BSDF_PRINCIPLED

            def get_flag(name):  # principledhelpers.h:95-106
                if name not in inner:
                    return False
                v = inner[name]
                return not (isinstance(v, float) and v == 0.0)
            has_trans = get_flag("spec_trans")
            if has_trans and flags & capi.BSDF_FLAG_TWOSIDED:
                raise RuntimeError("twosided: only materials without a transmission component can be nested!")   # twosided.cpp:52-57
            b.tex[0] = self.texture(inner, "base_color", 0.5)
            b.f[1] = self.scalar(inner, "roughness", 0.5)
            slots = {"anisotropic": 2, "metallic": 3, "clearcoat": 4, "sheen": 6, "sheen_tint": 7, "flatness": 8, "spec_tint": 9,
                     "spec_trans": 10}
            flags = 0
            for name, slot in slots.items():
                b.f[slot] = self.scalar(inner, name, 0.0)
                if get_flag(name):
                    flags |= capi.PRINCIPLED_HAS[name]
            b.f[5] = self.scalar(inner, "clearcoat_gloss", 0.0)
            b.u[0] = flags
            rates = [F32(inner.get("main_specular_sampling_rate", 1.0)), F32(inner.get("clearcoat_sampling_rate", 1.0)),
                     F32(inner.get("diffuse_reflectance_sampling_rate", 1.0))]
            b.u[1], b.u[2], b.u[3] = (int(np.asarray(r, dtype=F32).view(np.uint32)) for r in rates)
            if "eta" in inner and "specular" in inner:
                raise RuntimeError('Specified an invalid index of refraction property, either use "eta" or "specular" !')
            if "eta" in inner:
                b.f[0] = float(inner["eta"])
                if has_trans and b.f[0] == 1.0:       # "m_eta = 1 is not plausible for transmission" (principled.cpp:222-223)
                    b.f[0] = 1.001
            else:
                specular = F32(inner.get("specular", 0.5))
                if has_trans and specular == 0:        # principled.cpp:227-228
                    specular = F32(1e-3)
                b.f[0] = F32(2.0) * (F32(1.0) / (F32(1.0) - np.sqrt(F32(0.08) * specular, dtype=F32))) - F32(1.0)
        else:
            raise RuntimeError(f'Plugin "{t}" not found (bsdf outside the plt hot path)')
        self.bsdfs.append(b)
        self._bsdf_cache[key] = len(self.bsdfs) - 1
        return self._bsdf_cache[key]

    # ---- shapes + area emitters ----------------------------------------------------------------------------------------
    def shape(self, p):
        s = capi.Shape()
        s.bsdf, s.emitter = -1, -1
        t = p.plugin_name
        to_world = np.asarray(p.get("to_world", np.eye(4)), dtype=np.float64)
        bsdf_p = emitter_p = None
        for name, c in p.all_children():
            if c.cls == "bsdf":
                bsdf_p = c
            elif c.cls == "emitter":
                emitter_p = c
        s.bsdf = self.bsdf(bsdf_p if bsdf_p is not None else _DEFAULT_DIFFUSE)
        flip = bool(p.get("flip_normals", False))
        s.flags = capi.SHAPE_FLIP_NORMALS if flip else 0
        s.to_world, s.to_object = _m16(np.eye(4)), _m16(np.eye(4))
        idx = len(self.shapes)
        if t in ("ply", "serialized", "cube"):
            if t != "cube":
                path = p["filename"]
                if not os.path.isabs(path):
                    path = os.path.join(getattr(p, "base_path", "."), path)
                pos, nrm, uv, faces = load_ply(path) if t == "ply" else load_serialized(path, int(p.get("shape_index", 0)))
            else:
                pos, nrm, uv, faces = _cube_mesh()
            face_normals = bool(p.get("face_normals", False))
            pos = _xf_points(to_world, pos)
            if nrm is not None and not face_normals:
                nrm = _xf_normals(to_world, nrm)
                s.flags |= capi.SHAPE_HAS_NORMALS
            else:
                nrm = None
            if uv is not None:
                if p.get("flip_tex_coords", False):
                    uv = uv.copy(); uv[:, 1] = 1 - uv[:, 1]
                s.flags |= capi.SHAPE_HAS_UV
            if nrm is None and not face_normals:
                raise RuntimeError("ply without vertex normals: recompute_vertex_normals is not implemented; "
                                   "set face_normals=true")
            nv = len(pos)
            s.type = capi.SHAPE_MESH
            s.first_face, s.n_faces = sum(len(f) for f in self.faces), len(faces)
            self.positions.append(pos)
            self.normals.append(nrm if nrm is not None else np.zeros((nv, 3), F32))
            self.texcoords.append(uv if uv is not None else np.zeros((nv, 2), F32))
            self.faces.append(faces + np.uint32(self.n_vertices))
            self.face_shape.append(np.full(len(faces), idx, np.uint32))
            self.n_vertices += nv
            # Mesh::build_pmf (src/render/mesh.cpp:431-476, JIT branch): float32 areas, prefix sum
            p0, p1, p2 = pos[faces[:, 0]], pos[faces[:, 1]], pos[faces[:, 2]]
            area = (F32(.5) * np.linalg.norm(np.cross(p1 - p0, p2 - p0).astype(F32), axis=1)).astype(F32)
            cdf = np.cumsum(area, dtype=F32)
            s.cdf_offset = self.n_cdf
            self.area_cdf.append(cdf); self.n_cdf += len(cdf)
            s.area_sum = cdf[-1]; s.area_norm = F32(1) / cdf[-1]
            self._expand(pos)
        elif t == "rectangle":  # src/shapes/rectangle.cpp:95-140
            s.type = capi.SHAPE_RECTANGLE
            s.to_world, s.to_object = _m16(to_world), _m16(np.linalg.inv(to_world))
            du, dv = to_world[:3, 0] * 2, to_world[:3, 1] * 2
            n = np.linalg.inv(to_world[:3, :3]).T @ np.array([0, 0, 1.0])
            n /= np.linalg.norm(n)
            s.frame_n[:] = list(n.astype(F32))
            s.inv_area = F32(1.0 / np.linalg.norm(np.cross(du, dv)))
            s.flags |= capi.SHAPE_HAS_UV
            self._expand(_xf_points(to_world, np.array([[-1, -1, 0], [1, -1, 0], [1, 1, 0], [-1, 1, 0]], F32)))
        elif t == "sphere":  # src/shapes/sphere.cpp:120-175
            s.type = capi.SHAPE_SPHERE
            center = np.asarray(p.get("center", [0, 0, 0]), dtype=np.float64)
            radius = float(p.get("radius", 1.0))
            from .xmlscene import scale as _sc, translate as _tr
            tw = to_world @ _tr(center) @ _sc(radius)
            radius_w = float(np.linalg.norm(tw[:3, 0]))
            if abs(np.linalg.norm(tw[:3, 1]) - radius_w) > 1e-4 * radius_w:
                raise RuntimeError("'to_world' transform shouldn't contain any non-uniform scaling!")
            s.to_world, s.to_object = _m16(tw), _m16(np.linalg.inv(tw))
            s.center[:] = list(tw[:3, 3].astype(F32))
            s.radius = radius_w
            s.flags |= capi.SHAPE_HAS_UV
            c = tw[:3, 3]
            self._expand(np.array([c - radius_w, c + radius_w], F32))
        else:
            raise RuntimeError(f'Plugin "{t}" not found (shape outside the plt hot path)')
        self.shapes.append(s)
        if emitter_p is not None:
            if emitter_p.plugin_name != "area":
                raise RuntimeError("only 'area' emitters can be attached to shapes")
            e = capi.Emitter()
            e.type, e.shape = capi.EMITTER_AREA, idx
            e.radiance = self.texture(emitter_p, "radiance", None, within_emitter=True)
            if e.radiance < 0:  # default: d65 illuminant (src/emitters/area.cpp:62, Texture::D65(1.f))
                e.radiance = self.texture_plugin(Properties("d65"), True)
            if self.textures[e.radiance].type == capi.TEX_BITMAP:
                raise RuntimeError("spatially varying area lights are not supported on the plt path")
            e.sampling_weight = float(emitter_p.get("sampling_weight", 1.0))
            self.emitters.append(e)
            s.emitter = len(self.emitters) - 1
        return idx

    def _expand(self, pts):
        self.bbox[0] = np.minimum(self.bbox[0], pts.min(0)); self.bbox[1] = np.maximum(self.bbox[1], pts.max(0))

    # ---- standalone emitters -------------------------------------------------------------------------------------------
    def emitter(self, p):
        t = p.plugin_name
        if t == "constant":  # src/emitters/constant.cpp:52-67: props.texture_d65("radiance", 1.f)
            if any(x.type in (capi.EMITTER_CONSTANT, capi.EMITTER_ENVMAP) for x in self.emitters):
                raise RuntimeError("Only one environment emitter can be specified per scene.")   # scene.cpp:60-67
            e = capi.Emitter()
            e.type, e.shape = capi.EMITTER_CONSTANT, -1
            v = p.get("radiance")
            if v is None or (isinstance(v, (int, float)) and not isinstance(v, bool)):
                e.radiance = self.texture_plugin(Properties("d65", scale=1.0 if v is None else float(v)), True)
            else:
                e.radiance = self.texture(p, "radiance", None, within_emitter=True)
            if self.textures[e.radiance].type == capi.TEX_BITMAP:
                raise RuntimeError("Expected a non-spatially varying radiance spectra!")
            e.sampling_weight = float(p.get("sampling_weight", 1.0))
            self.emitters.append(e)
            return
        if t == "envmap":
            return self.envmap(p)
        if t not in ("directional", "directionalspot"):
            raise RuntimeError(f'Plugin "{t}" not found (emitter outside the plt hot path)')
        e = capi.Emitter()
        e.type, e.shape = capi.EMITTER_DIRECTIONAL, -1
        tw = np.asarray(getattr(p.get("to_world"), "matrix", p.get("to_world", np.eye(4))), dtype=np.float64).reshape(4, 4)
        if "direction" in p:  # src/emitters/directional.cpp:64-78
            if "to_world" in p:
                raise RuntimeError("Only one of the parameters 'direction' and 'to_world' can be specified at the same time!'")
            d = np.asarray(p["direction"], dtype=np.float64)
            if t == "directionalspot":   # look_at(0, direction, up) with up from coordinate_system(direction) (directionalspot.cpp:70-79)
                from .xmlscene import look_at
                dn = d / np.linalg.norm(d)
                sgn = 1.0 if dn[2] >= 0 else -1.0
                a = -1.0 / (sgn + dn[2]); b = dn[0] * dn[1] * a
                up = np.array([1.0 + sgn * dn[0] * dn[0] * a, sgn * b, -sgn * dn[0]])
                tw = look_at(np.zeros(3), dn, up)
        else:
            d = tw[:3, :3] @ np.array([0, 0, 1.0])
        d = d / np.linalg.norm(d)
        e.direction[:] = list(d.astype(F32))
        if t == "directionalspot":   # src/emitters/directionalspot.cpp:60-100: direction sampled in a cone of half-angle `spread_angle`
            e.type = capi.EMITTER_DIRECTIONALSPOT
            e.to_world[:] = list(tw.astype(F32).reshape(-1))
            e.to_local[:] = list(np.linalg.inv(tw).astype(F32).reshape(-1))
            e.env_pad[0] = float(np.sin(F32(p.get("spread_angle", 0.0)), dtype=F32))
        e.radiance = self.texture(p, "irradiance", None, within_emitter=True)
        if e.radiance < 0:
            e.radiance = self.texture_plugin(Properties("d65"), True)
        e.sampling_weight = float(p.get("sampling_weight", 1.0))
        self.emitters.append(e)

    def envmap(self, p):
        """EnvironmentMapEmitter (src/emitters/envmap.cpp:110-243): latitude-longitude bitmap -> per-texel rgb2spec coefficients
        + scale (extra column for the periodic boundary) and a Hierarchical2D over luminance * sin(theta)."""
        from . import distr2d, imageio
        if any(x.type in (capi.EMITTER_CONSTANT, capi.EMITTER_ENVMAP) for x in self.emitters):
            raise RuntimeError("Only one environment emitter can be specified per scene.")   # scene.cpp:60-67
        if "bitmap" in p:
            if "filename" in p:
                raise RuntimeError('Cannot specify both "bitmap" and "filename".')
            b = p["bitmap"]
            rgb = np.asarray(getattr(b, "array", b), dtype=F32)
        else:
            path = p["filename"]
            if not os.path.isabs(path):
                path = os.path.join(getattr(p, "base_path", "."), path)
            ext = path.rsplit(".", 1)[-1].lower()
            if ext == "exr":
                rgb = imageio.read_exr(path)[0]
            elif ext == "npy":
                rgb = np.load(path).astype(F32)
            else:   # 8-bit formats are sRGB-encoded: Bitmap::convert(..., srgb_gamma = false) linearises them
                from PIL import Image
                rgb = spectra.srgb_to_linear(np.asarray(Image.open(path).convert("RGB"), dtype=F32) / F32(255))
        if rgb.ndim == 2:
            rgb = rgb[..., None]
        if rgb.shape[2] == 1:
            rgb = np.repeat(rgb, 3, axis=2)
        rgb = np.ascontiguousarray(rgb[..., :3], dtype=F32)
        H, W = rgb.shape[:2]
        if W < 2 or H < 3:
            raise RuntimeError("the environment map resolution must be at least 2x3 pixels")
        lum = (F32(0.212671) * rgb[..., 0] + F32(0.715160) * rgb[..., 1] + F32(0.072169) * rgb[..., 2]).astype(F32)
        offset = F32(0)
        if p.get("mis_compensation", False):   # envmap.cpp:159-181
            offset = F32(lum.astype(np.float64).sum() / (W * H))
            if offset - min(F32(0), lum.min()) <= F32(0.01) * offset:
                offset = F32(0)
        if self.rgb:   # envmap.cpp:199-200: coeff = rgb (kept 4 floats wide; the fourth is unused)
            data = np.concatenate([rgb, np.zeros_like(rgb[..., :1])], axis=2).astype(F32)
        else:
            scale = (rgb.max(axis=2) * F32(2)).astype(F32)
            coeff = spectra.srgb_model_fetch(rgb / np.maximum(F32(1e-8), scale)[..., None])
            data = np.concatenate([coeff, scale[..., None]], axis=2).astype(F32)
        data = np.concatenate([data, data[:, :1]], axis=1)                    # last column mirrors the first
        theta_scale = F32(1.0) / F32(H - 1) * F32(np.pi)
        sin_theta = np.sin((np.arange(H, dtype=F32) * theta_scale).astype(F32)).astype(F32)
        lw = (np.maximum(lum - offset, F32(0)) * sin_theta[:, None]).astype(F32)
        lw = np.concatenate([lw, lw[:, :1]], axis=1)
        e = capi.Emitter()
        e.type, e.shape = capi.EMITTER_ENVMAP, -1
        e.env_res[:] = [W + 1, H]
        e.env_scale = float(p.get("scale", 1.0))
        e.env_warp = self._push_data(distr2d.flatten_hierarchical(distr2d.build_hierarchical(lw), (W + 1, H)))
        e.env_data = self._push_data(data.reshape(-1))
        e.radiance = self.texture_plugin(Properties("d65"), False)             # m_d65 = Texture::D65(1.f)
        tw = np.asarray(getattr(p.get("to_world"), "matrix", p.get("to_world", np.eye(4))), dtype=np.float64).reshape(4, 4)
        e.to_world[:] = list(tw.astype(F32).reshape(-1))
        e.to_local[:] = list(np.linalg.inv(tw).astype(F32).reshape(-1))
        e.sampling_weight = float(p.get("sampling_weight", 1.0))
        self.emitters.append(e)

    # ---- sensors / film --------------------------------------------------------------------------------------------------
    def sensor(self, p, film_size=None):
        t = p.plugin_name
        if t == "batch":  # src/sensors/batch.cpp:70-128
            subs = [c for _, c in p.all_children() if c.cls == "sensor"]
            if not subs:
                raise RuntimeError("BatchSensor: at least one child sensor must be specified!")
            self._film_and_sampler(p)
            w = self.film.width
            if w % len(subs):
                raise RuntimeError("BatchSensor: the horizontal resolution must be divisible by the number of sensors")
            self.film.width = w // len(subs)
            for c in subs:
                self.sensor(c, film_size=(self.film.width, self.film.height))
            return
        if t not in ("perspective", "orthographic"):
            raise RuntimeError(f'Plugin "{t}" not found (sensor outside the plt hot path)')
        if film_size is None:
            self._film_and_sampler(p)
            film_size = (self.film.width, self.film.height)
        s = capi.Sensor()
        s.srf = -1
        if "srf" in p:
            if self.rgb:
                raise RuntimeError("sensor 'srf': spectral response functions need a spectral variant")
            s.srf = self.texture(p, "srf", None)
            if self.textures[s.srf].type not in (capi.TEX_REGULAR, capi.TEX_IRREGULAR):
                raise RuntimeError("sensor 'srf' must be a regular/irregular spectrum")
        near, far = float(p.get("near_clip", 1e-2)), float(p.get("far_clip", 1e4))
        to_world = np.asarray(p.get("to_world", np.eye(4)), dtype=np.float64)
        W, H = film_size
        aspect = W / H
        from .xmlscene import scale as _sc, translate as _tr
        if t == "perspective":
            s.type = capi.SENSOR_PERSPECTIVE
            x_fov = _parse_fov(p, aspect)
            # Transform::perspective (include/mitsuba/core/transform.h) + perspective_projection (sensor.h:230-265)
            recip = 1.0 / (far - near)
            cot = 1.0 / math.tan(math.radians(x_fov * 0.5))
            proj = np.array([[cot, 0, 0, 0], [0, cot, 0, 0], [0, 0, far * recip, -near * far * recip], [0, 0, 1, 0]])
        else:
            s.type = capi.SENSOR_ORTHOGRAPHIC
            proj = _sc([1, 1, 1.0 / (far - near)]) @ _tr([0, 0, -near])
        cam_to_sample = _sc([-0.5, -0.5 * aspect, 1]) @ _tr([-1, -1 / aspect, 0]) @ proj
        s.sample_to_camera = _m16(np.linalg.inv(cam_to_sample))
        s.to_world = _m16(to_world)
        s.near_clip, s.far_clip = near, far
        s.pp_offset[0] = float(p.get("principal_point_offset_x", 0.0))
        s.pp_offset[1] = float(p.get("principal_point_offset_y", 0.0))
        self.sensors.append(s)
        self.sensor_ids.append(p.id)

    def _film_and_sampler(self, p):
        film_p = sampler_p = None
        for _, c in p.all_children():
            if c.cls == "film":
                film_p = c
            elif c.cls == "sampler":
                sampler_p = c
        f = capi.Film()
        film_p = film_p if film_p is not None else Properties("hdrfilm")
        if film_p.plugin_name != "hdrfilm":
            raise RuntimeError(f'Plugin "{film_p.plugin_name}" not found (film outside the plt hot path)')
        f.width, f.height = int(film_p.get("width", 768)), int(film_p.get("height", 576))
        if any(k in film_p for k in ("crop_width", "crop_height", "crop_offset_x", "crop_offset_y")):
            raise RuntimeError("hdrfilm: crop windows are not supported")
        rf = None
        for _, c in film_p.all_children():
            if c.cls == "rfilter":
                rf = c
        name = rf.plugin_name if rf is not None else "gaussian"
        if name == "gaussian":  # src/rfilters/gaussian.cpp:48-53
            f.rfilter = capi.RFILTER_GAUSSIAN
            f.stddev = float(rf.get("stddev", 0.5)) if rf is not None else 0.5
            f.radius = 4 * f.stddev
        elif name == "tent":
            f.rfilter = capi.RFILTER_TENT
            f.radius = float(rf.get("radius", 1.0))
        elif name == "box":
            f.rfilter = capi.RFILTER_BOX
            f.radius = 0.5
        else:
            raise RuntimeError(f'Plugin "{name}" not found (rfilter outside the plt hot path)')
        self.film = f
        self.pixel_format = film_p.get("pixel_format", "rgb")
        if sampler_p is not None:
            if sampler_p.plugin_name != "independent":
                raise RuntimeError("only the 'independent' sampler is supported on the plt path")
            self.spp = int(sampler_p.get("sample_count", 4))
            self.base_seed = int(sampler_p.get("seed", 0))

    base_seed = 0

    # ---- finalize ----------------------------------------------------------------------------------------------------------
    def finalize(self):
        def cat(lst, shape, dt):
            return np.ascontiguousarray(np.concatenate(lst).astype(dt)) if lst else np.zeros(shape, dt)
        self.a_pos = cat(self.positions, (0, 3), F32)
        self.a_nrm = cat(self.normals, (0, 3), F32)
        self.a_uv = cat(self.texcoords, (0, 2), F32)
        self.a_faces = cat(self.faces, (0, 3), np.uint32)
        self.a_face_shape = cat(self.face_shape, (0,), np.uint32)
        self.a_tex_data = cat(self.tex_data, (0,), F32)
        if self.a_tex_data.size == 0:
            self.a_tex_data = np.zeros(1, F32)
        self.a_cdf = cat(self.area_cdf, (0,), F32)
        if self.a_cdf.size == 0:
            self.a_cdf = np.zeros(1, F32)
        tabs = spectra.cie_tables()
        self.a_cie = np.ascontiguousarray(np.concatenate([tabs["srgb"], tabs["xyz"][1:2], tabs["xyz"][0:1], tabs["xyz"][2:3]]),
                                          dtype=F32)   # R, G, B, Y, X, Z rows
        # scene bounding sphere for directional emitters (src/render/scene.cpp:70-80, BoundingSphere from bbox)
        if np.all(np.isfinite(self.bbox[0])):
            c = (self.bbox[0] + self.bbox[1]) * 0.5
            r = float(np.linalg.norm(self.bbox[1] - c))
        else:
            c, r = np.zeros(3), 0.0
        for e in self.emitters:
            if e.type == capi.EMITTER_DIRECTIONAL:
                e.bsphere_center[:] = list(c.astype(F32))
                e.bsphere_radius = r
            elif e.type in (capi.EMITTER_CONSTANT, capi.EMITTER_ENVMAP, capi.EMITTER_DIRECTIONALSPOT):
                # ::set_scene (constant.cpp:74-89, envmap.cpp:327-341, directionalspot.cpp:102-112)
                ray_eps = F32(2.0 ** -24 * 1500)   # math::RayEpsilon = dr::Epsilon<float> * 1500 (include/mitsuba/core/math.h:18)
                e.bsphere_center[:] = list(c.astype(F32))
                e.bsphere_radius = max(ray_eps, F32(r) * (F32(1) + ray_eps)) if np.all(np.isfinite(self.bbox[0])) else ray_eps
        self.c_shapes = (capi.Shape * max(1, len(self.shapes)))(*self.shapes)
        self.c_bsdfs = (capi.Bsdf * max(1, len(self.bsdfs)))(*self.bsdfs)
        self.c_textures = (capi.Texture * max(1, len(self.textures)))(*self.textures)
        self.c_emitters = (capi.Emitter * max(1, len(self.emitters)))(*self.emitters)
        self.c_sensors = (capi.Sensor * max(1, len(self.sensors)))(*self.sensors)
        d = capi.SceneDesc()
        d.abi_version = capi.PLTB_ABI_VERSION
        d.n_vertices, d.n_faces, d.n_shapes = len(self.a_pos), len(self.a_faces), len(self.shapes)
        d.n_bsdfs, d.n_textures, d.n_emitters = len(self.bsdfs), len(self.textures), len(self.emitters)
        d.n_sensors = len(self.sensors)
        d.n_tex_data, d.n_area_cdf = self.a_tex_data.size, self.a_cdf.size
        d.positions, d.normals, d.texcoords = capi.fptr(self.a_pos), capi.fptr(self.a_nrm), capi.fptr(self.a_uv)
        d.faces, d.face_shape = capi.uptr(self.a_faces), capi.uptr(self.a_face_shape)
        d.shapes, d.bsdfs, d.textures = self.c_shapes, self.c_bsdfs, self.c_textures
        d.tex_data, d.emitters, d.area_cdf = capi.fptr(self.a_tex_data), self.c_emitters, capi.fptr(self.a_cdf)
        d.sensors = self.c_sensors
        d.film = self.film
        d.cie_srgb = capi.fptr(self.a_cie)
        self.desc = d
        return d


_DEFAULT_DIFFUSE = Properties("diffuse")
_DEFAULT_DIFFUSE.cls = "bsdf"

_IOR = {"vacuum": 1.0, "air": 1.00028, "water": 1.3330, "bk7": 1.5046, "diamond": 2.419, "acrylic glass": 1.49,
        "polypropylene": 1.49, "pyrex": 1.470, "sodium chloride": 1.544, "amber": 1.55, "pet": 1.5750,
        "water ice": 1.31, "fused quartz": 1.458, "acetone": 1.36, "ethanol": 1.361, "helium": 1.00004,
        "hydrogen": 1.00013, "carbon dioxide": 1.00045, "carbon tetrachloride": 1.461, "glycerol": 1.4729,
        "benzene": 1.501, "silicone oil": 1.52045, "bromine": 1.661, "silicon": 3.4777}


def _lookup_ior(v):
    """include/mitsuba/render/ior.h"""
    if isinstance(v, str):
        if v.lower() not in _IOR:
            raise RuntimeError(f'Unable to find an IOR value for "{v}"')
        return _IOR[v.lower()]
    return float(v)


def _parse_fov(p, aspect):
    """src/render/sensor.cpp parse_fov (fov + fov_axis form)"""
    if "focal_length" in p and "fov" in p:
        raise RuntimeError("Please specify either a focal length ('focal_length') or a field of view ('fov')!")
    if "fov" not in p:
        f = str(p.get("focal_length", "50mm")).replace("mm", "")
        fov = 2 * math.degrees(math.atan(math.sqrt(36 ** 2 + 24 ** 2) / (2 * float(f))))
        axis = "diagonal"
    else:
        fov = float(p["fov"])
        if fov <= 0 or fov >= 180:
            raise RuntimeError("The horizontal field of view must be in the range [0, 180]!")
        axis = str(p.get("fov_axis", "x")).lower()
    if axis == "smaller":
        axis = "x" if aspect < 1 else "y"
    elif axis == "larger":
        axis = "y" if aspect < 1 else "x"
    if axis == "x":
        return fov
    if axis == "y":
        return math.degrees(2 * math.atan(math.tan(0.5 * math.radians(fov)) * aspect))
    if axis == "diagonal":
        diag = 2 * math.tan(0.5 * math.radians(fov))
        width = diag / math.sqrt(1 + 1 / (aspect * aspect))
        return math.degrees(2 * math.atan(width * 0.5))
    raise RuntimeError("The 'fov_axis' parameter must be set to one of 'smaller', 'larger', 'diagonal', 'x', or 'y'!")


def flatten(tree, rgb=False):
    """Scene ctor (src/render/scene.cpp:22-96): walk the children of the <scene> node.  rgb: instantiate the plugins as an RGB
    variant does (is_rgb_v<Spectrum>)."""
    fs = FlatScene(rgb)
    if tree.cls != "scene":
        raise RuntimeError("flatten(): expected a scene node")
    for _, c in tree.all_children():
        if c.cls == "shape":
            fs.shape(c)
        elif c.cls == "emitter":
            fs.emitter(c)
        elif c.cls == "sensor":
            if fs.sensors:
                raise RuntimeError("multiple top-level sensors: use a 'batch' sensor")
            fs.sensor(c)
        elif c.cls == "integrator":
            fs.integrator_props = c
        elif c.cls in ("bsdf", "texture", "spectrum"):
            pass  # referenced by id
        else:
            raise RuntimeError(f"<{c.cls}> is not supported at the top level of a scene")
    if not fs.sensors:
        raise RuntimeError("scene has no sensor")
    fs.finalize()
    return fs
