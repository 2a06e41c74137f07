"""The PLT value types the reference binds to Python (src/plt/python/plt_v.cpp:6-77 and the header types behind them):
`DiffractionGratingType`, `DiffractionGrating3f`, `Coherence3f`, `GeneralizedRadiance3f`, `PLTInteraction3f`,
`PLTSamplePhaseData3f`, `BounceData3f`, `PLTBeam3f`, plus the scalar helpers the reference's scene-building scripts use next
to them (`Frame3f`, `ScalarVector2f/2i/3f`, `spectrum_to_file` / `spectrum_from_file`).

These are host-side, scalar (float32 numpy) mirrors with the reference's constructor signatures, field names and methods, so
that `scripts/spectrograph/scene.py:57-144` builds its scene against this module with only the import changed.  The render path
does not go through them: coherence is inert in the reference's integrator and the kernels carry the grating
parameters in `pltb_bsdf`.  `DiffractionGrating3f` follows include/mitsuba/plt/diffractiongrating.h:32-290 operation by
operation and is pinned on the same known answers as the oracle (tests/golden/ref_kat.json).
"""
import math

import numpy as np

F = np.float32


def ScalarVector2f(*a):
    return np.array(a[0] if len(a) == 1 and np.ndim(a[0]) else (a * 2 if len(a) == 1 else a), F).reshape(2)


def ScalarVector3f(*a):
    return np.array(a[0] if len(a) == 1 and np.ndim(a[0]) else (a * 3 if len(a) == 1 else a), F).reshape(3)


def ScalarVector2i(*a):
    return np.array(a[0] if len(a) == 1 and np.ndim(a[0]) else (a * 2 if len(a) == 1 else a), np.int32).reshape(2)


def _fma(a, b, c):
    return F(np.float64(a) * np.float64(b) + np.float64(c))   # exact product and sum in float64, one rounding to float32


class Frame3f:
    """mi.Frame3f(s, t, n) / mi.Frame3f(n) (include/mitsuba/core/frame.h)"""

    def __init__(self, *a):
        if len(a) == 1:
            self.n = ScalarVector3f(a[0])
            self.s, self.t = coordinate_system(self.n)
        else:
            self.s, self.t, self.n = (ScalarVector3f(x) for x in a)

    def to_local(self, v):
        v = ScalarVector3f(np.ravel(v))
        return np.array([np.dot(v, self.s), np.dot(v, self.t), np.dot(v, self.n)], F)

    def to_world(self, v):
        v = ScalarVector3f(np.ravel(v))
        return (self.s * v[0] + self.t * v[1] + self.n * v[2]).astype(F)


def coordinate_system(n):
    """include/mitsuba/core/vector.h:116-136"""
    n = ScalarVector3f(n)
    sign = F(math.copysign(1.0, n[2]))
    a = F(-1) / (sign + n[2])
    b = n[0] * n[1] * a
    s = np.array([F(n[0] * n[0] * a * sign) + F(1), F(b * sign), F(-n[0] * sign)], F)   # mulsign(x, n.z) == x * sign
    t = np.array([b, _fma(n[1], n[1] * a, sign), -n[1]], F)
    return s, t


class DiffractionGratingType:
    """include/mitsuba/plt/diffractiongrating.h:13-18 (flag values)"""
    Sinusoidal, Rectangular, Linear, Radial = 0x00, 0x01, 0x02, 0x10


def _bessel_j(x, nu, points=20, c=45.0, f=1.5):
    """math::bessel_j (include/mitsuba/core/math.h:204-223, 295-347) in float32"""
    x = F(x)
    nua = abs(int(nu))
    h = F(F(np.pi) / F(points - 1))
    t = (F(0) + h * np.arange(points, dtype=F)).astype(F)
    t[-1] = F(np.pi)
    g = np.cos(F(nua) * t - x * np.sin(t).astype(F)).astype(F)
    small = F(1 / np.pi) * (h / F(3) * (g[0] + F(4) * g[1:-1:2].sum(dtype=F) + F(2) * g[2:-1:2].sum(dtype=F) + g[-1]))
    ax = abs(x)
    if ax > 10 * np.finfo(F).eps / 2:
        sign = F(1) if (x >= 0 or nua % 2) else F(-1)
        asymp = sign * F(math.sqrt(2.0 / (math.pi * ax))) * F(math.cos(ax - math.pi * nua / 2 - math.pi / 4))
    else:
        asymp = F(1 if nua == 0 else 0)
    w = F(1.0 / (1.0 + math.exp(-f * (ax - (c + nua)))))
    return F((-1 if nu < 0 else 1) * ((F(1) - w) * small + w * asymp))


class DiffractionGrating3f:
    """DiffractionGrating<Float, Spectrum> (include/mitsuba/plt/diffractiongrating.h:32-290):
    DiffractionGrating3f(grating_angle, inv_period, q, lobes, type, multiplier, uv)"""

    def __init__(self, grating_angle, inv_period, q, lobes, type, multiplier, uv):
        self.inv_period = ScalarVector2f(inv_period)
        self.q, self.lobes, self.type, self.multiplier = F(q), int(lobes), int(type), F(multiplier)
        ang = F(grating_angle)
        if (self.type & DiffractionGratingType.Radial) == 0:
            self.grating_dir = np.array([math.cos(ang), math.sin(ang)], F)
        else:
            r = ScalarVector2f(uv) - F(0.5)
            r = r / F(math.sqrt(r[0] * r[0] + r[1] * r[1]))
            c, s = F(math.cos(ang)), F(math.sin(ang))
            vx, vy = r[0], -r[1]
            self.grating_dir = np.array([_fma(s, vy, c * vx), _fma(c, vy, -s * vx)], F)

    def is_1D_grating(self):
        return bool(self.inv_period[1] < np.finfo(F).eps / 2)

    def alpha(self, wi, k):
        wi = ScalarVector3f(np.ravel(wi))
        return F(math.exp(-float(F(wi[2] * self.q * F(k)) ** 2)))

    def lobe_intensity(self, lobe, wi, wl):
        lobe, wi = ScalarVector2i(np.ravel(lobe)), ScalarVector3f(np.ravel(wi))
        a = F(F(4 * np.pi) * self.q / (F(wl) * abs(wi[2])))
        gtype = self.type & 0xF

        def axis(l):
            if l == 0:
                return F(1)
            if gtype == DiffractionGratingType.Sinusoidal:
                j = _bessel_j(a, int(l))
                return F(j * j)
            if gtype == DiffractionGratingType.Rectangular:
                x = F(np.pi) * F(l) / F(2)
                return F(math.sin(a / F(2))) * F(math.sin(x) / x)
            return F(1) / F(math.sqrt(abs(float(l))))
        ix = axis(lobe[0])
        iy = ix if self.is_1D_grating() else axis(lobe[1])
        return F(self.multiplier * ix * iy)

    def sample_lobe(self, sample2, wi, wl):
        n = self.lobes // 2 + 1
        inten = []
        for l in range(n):
            li = self.lobe_intensity((l, 0), wi, wl)
            inten.append(F(li / F(2)) if l == 0 else li)
        total = F(0)
        for v in inten:
            total = F(total + v)
        rn = (ScalarVector2f(sample2) - F(0.5)) * F(2)
        cdf, lobe, pdf = F(0), [F(0), F(0)], [F(0), F(0)]
        for l in range(n):
            p = F(inten[l] / total)
            for ax in (0, 1):
                if abs(rn[ax]) > cdf:
                    pdf[ax] = p if l == 0 else F(p / F(2))
                    lobe[ax] = F(l)
            cdf = F(cdf + p)
        sgn = [F(math.copysign(1.0, rn[0])), F(math.copysign(1.0, rn[1]))]
        return ScalarVector2i(int(lobe[0] * sgn[0]), int(lobe[1] * sgn[1])), ScalarVector2f(pdf[0], pdf[1])

    def lobe_pdf(self, lobe, wi, wl):
        lobe = ScalarVector2i(np.ravel(lobe))
        total, ix, iy = F(0), F(0), F(0)
        for l in range(self.lobes // 2 + 1):
            li = self.lobe_intensity((l, 0), wi, wl)
            if l == 0:
                li = F(li * F(0.5))
            total = F(total + li)
            if abs(int(lobe[0])) == l:
                ix = li
            if abs(int(lobe[1])) == l:
                iy = li
        return F(F(ix / total) * F(iy / total))

    def diffract(self, wi, lobe, wl):
        """-> (direction, valid)"""
        wi, lobe, wl = ScalarVector3f(np.ravel(wi)), ScalarVector2i(np.ravel(lobe)), F(wl)
        eps = np.finfo(F).eps / 2
        px, py = F(math.sqrt(F(wi[0] * wi[0]) + F(wi[2] * wi[2]))), F(math.sqrt(F(wi[1] * wi[1]) + F(wi[2] * wi[2])))
        sin_ix = F(wi[0] / px) if px > eps else F(0)
        sin_iy = F(wi[1] / py) if py > eps else F(0)
        cg, sg = self.grating_dir
        lrx = F(F(cg * F(lobe[0])) - F(sg * F(lobe[1])))
        lry = F(F(sg * F(lobe[0])) + F(cg * F(lobe[1])))
        a = F(F(F(wl * lrx) * self.inv_period[0]) - sin_ix)
        b = F(F(F(wl * lry) * self.inv_period[1]) - sin_iy)
        with np.errstate(all="ignore"):
            m = F(F(F(a * a) - F(1)) / F(F(F(a * b) * F(a * b)) - F(1)))
            qq = F(F(1) - F(F(b * b) * m))
            d = np.array([a * F(np.sqrt(max(F(0), qq))), b * F(np.sqrt(m)),
                          F(np.sqrt(max(F(0), F(F(F(1) - F(F(a * a) * qq)) - F(F(b * b) * m)))))], F)
        return d, bool(abs(a) <= 1 and abs(b) <= 1)


class Coherence3f:
    """Coherence<Float, Spectrum> (include/mitsuba/plt/plt.h:22-100): Coherence3f(diffusivity | 2x2 matrix, l)"""

    def __init__(self, diff=0.0, l=0.0):
        if isinstance(diff, Coherence3f):
            self.dmat, self.opl = diff.dmat.copy(), F(diff.opl)
        else:
            self.dmat = np.array(diff, F).reshape(2, 2) if np.ndim(diff) else np.diag([F(diff), F(diff)]).astype(F)
            self.opl = F(l)

    def rmm(self):
        return F(self.opl * F(1e3))

    def propagate(self, rd, mask=True):
        if mask:
            self.opl = F(self.opl + F(rd))

    def inv_coherence_matrix(self, k=None):
        s = F(1) / self.rmm() if k is None else F(k) / (F(2 * np.pi) * self.rmm())
        return (s * self.dmat).astype(F)

    def inv_coherence_det(self, k=None):
        m = self.inv_coherence_matrix(k)
        return F(_fma(m[0, 0], m[1, 1], -F(m[0, 1] * m[1, 0])))

    def transform(self, U, mask=True):
        if mask:
            U = np.array(U, F).reshape(2, 2)
            self.dmat = (U.T @ (self.dmat @ U)).astype(F)

    def __repr__(self):
        return f"Coherence[dmat = {self.dmat.tolist()}, opl = {float(self.opl)}]"


class _Struct:
    _fields = ()

    def __init__(self, *a, **kw):
        for k in self._fields:
            setattr(self, k, None)
        for k, v in zip(self._fields, a):
            setattr(self, k, v)
        for k, v in kw.items():
            if k not in self._fields:
                raise TypeError(f"{type(self).__name__}: unexpected field '{k}'")
            setattr(self, k, v)

    def __repr__(self):
        return f"{type(self).__name__}[" + ", ".join(f"{k} = {getattr(self, k)!r}" for k in self._fields) + "]"


class GeneralizedRadiance3f(_Struct):
    """include/mitsuba/plt/plt.h:146-171: only L is ever filled; L1..L3 = 0, coherence = Coherence(1e-3, 0)"""
    _fields = ("L", "L1", "L2", "L3", "coherence")

    def __init__(self, L=0.0):
        super().__init__(L, 0.0, 0.0, 0.0, Coherence3f(1e-3, 0.0))


class PLTInteraction3f(_Struct):
    """include/mitsuba/plt/plt.h:103-140"""
    _fields = ("sp_coherence", "t")


class PLTSamplePhaseData3f(_Struct):
    """include/mitsuba/plt/sample_solve.h:10-54: PLTSamplePhaseData3f(bsdf_sample, diffraction_lobe, sampled_normal, coherence,
    sampling_wavelengths)"""
    _fields = ("bsdf_sample", "diffraction_lobe", "sampled_normal", "coherence", "sampling_wavelengths")


class BounceData3f(_Struct):
    """include/mitsuba/plt/bouncebuffer.h:21-95: the record plt_sample_phase writes per vertex (plt.py:145-160)"""
    _fields = ("id", "interaction", "wi", "wo", "bsdf_flags", "rr_thp", "throughput", "bsdf_weight", "is_emitter", "last_nd_pdf",
               "sampled_lobe", "sampling_wavelengths", "active")


class PLTBeam3f:
    """PLTBeam<Float, Spectrum> (include/mitsuba/plt/beam.h:16-208): bound to Python, never used by plt.py (its sourcing /
    measurement steps are TODOs, plt.py:302-311,475-490).  PLTBeam3f(sp, diff, origin, dir, tangent, distant, active=True)"""

    def __init__(self, sp, diff, origin, dir, tangent, distant, active=True):
        self.sp, self.phase = sp, F(0)
        self.origin, self.dir, self.tangent = ScalarVector3f(origin), ScalarVector3f(dir), ScalarVector3f(tangent)
        self.distant, self.active = bool(distant), bool(active)
        self.coherence = Coherence3f(np.array(diff, F).reshape(2, 2), 0.001 if distant else 0.0)

    def scale_sp(self, s):
        self.sp = self.sp * s

    def transverse_rotation(self):
        vert = np.cross(self.tangent, self.dir).astype(F)
        return np.array([[self.tangent[0], self.tangent[1]], [vert[0], vert[1]]], F).T.copy()

    def inv_coherence(self, k=None):
        return self.coherence.inv_coherence_matrix(k)

    def mutual_coherence(self, k, diff):
        dxy = ScalarVector3f(np.ravel(diff))[:2]
        e = F(np.dot(dxy, self.inv_coherence() @ dxy))
        if np.ndim(k):   # Wavelength overload (spectral variants): exp(-0.5 * k / (2 pi) * dxy^T C^-1 dxy) per wavelength
            return np.exp(F(-0.5) * (np.asarray(k, F) * F(1 / (2 * np.pi))) * e).astype(F)
        return F(math.exp(-0.5 * float(e)))

    def propagate(self, t, mask=True):
        self.coherence.propagate(t, mask)
        if mask:
            self.origin = (self.origin + self.dir * F(t)).astype(F)


def spectrum_to_file(path, wavelengths, values):
    """mi.spectrum_to_file (include/mitsuba/core/spectrum.h:527-534, src/core/spectrum.cpp): one "wavelength value" pair per line"""
    wavelengths, values = np.ravel(np.asarray(wavelengths, np.float64)), np.ravel(np.asarray(values, np.float64))
    if len(wavelengths) != len(values):
        raise RuntimeError("spectrum_to_file(): wavelengths and values must have the same length")
    with open(path, "w") as f:
        for w, v in zip(wavelengths, values):
            f.write(f"{float(F(w))!r} {float(F(v))!r}\n")


def spectrum_from_file(path):
    """mi.spectrum_from_file (include/mitsuba/core/spectrum.h:515-525): -> (wavelengths, values)"""
    from .xmlscene import read_spd
    return read_spd(path)
