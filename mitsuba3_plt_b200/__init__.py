"""mitsuba3_plt_b200 -- B200-native drop-in for the forward render path of Mitsuba-PLT's `plt` integrator.

The module mirrors the slice of the `mitsuba` Python surface that the reference's PLT scripts use
(scripts/rendering/main-headless.py, scripts/rendering/utils/render.py, scripts/spectrograph/*.py):

    import mitsuba3_plt_b200 as mi
    mi.set_variant("cuda_spectral")
    scene = mi.load_file("scenes/cbox-grating/cbox.xml", resx=1024, resy=1024)
    integrator = mi.load_dict({"type": "plt", "max_depth": 7, "rr_depth": 50})
    image = mi.render(scene, integrator=integrator, spp=1024, seed=0)      # (H, W, 3) float32

Reference surfaces mirrored (paths relative to the reference tree):
  mi.render                      src/python/python/util.py:378-507
  Integrator.render              include/mitsuba/render/integrator.h:77, src/python/python/ad/integrators/common.py:46-110
  PLTIntegrator(props)           scripts/rendering/integrators/plt.py:13-48, registered at :531
  mi.load_file / mi.load_dict    src/core/python/xml_v.cpp:68-116
  mi.traverse / params.update    src/python/python/util.py:269
  mi.register_integrator         include/mitsuba/python/python.h:55-90

Everything per-sample runs in libpltb200.so (hand-written sm_100a CUDA behind the C ABI of include/pltb.h).  There is NO
CPU fallback: rendering without the built library or without a CUDA device raises RuntimeError.
"""
import ctypes as C
import os

import numpy as np

from . import capi, imageio
from .scene import FlatScene, flatten
from .xmlscene import Color, Properties, SpectrumValue, Transform, dict_to_tree, load_file_tree
from .plt_types import (BounceData3f, Coherence3f, DiffractionGrating3f, DiffractionGratingType, Frame3f,  # noqa: F401
                        GeneralizedRadiance3f, PLTBeam3f, PLTInteraction3f, PLTSamplePhaseData3f, ScalarVector2f, ScalarVector2i,
                        ScalarVector3f, spectrum_from_file, spectrum_to_file)

__all__ = ["set_variant", "variant", "load_file", "load_dict", "render", "traverse", "register_integrator",
           "PLTIntegrator", "Scene", "Properties", "ScalarTransform4f", "Transform4f", "Context", "context",
           "init_distributed", "Bitmap", "util", "is_spectral", "is_polarized", "is_rgb", "StokesIntegrator"]

ScalarTransform4f = Transform4f = Transform


class _Namespace:
    """`mi.scalar_rgb.X`: the reference's scene-building scripts reach the scalar types through a variant module
    (scripts/spectrograph/scene.py:75-86); here every variant exposes the same host-side types"""
    DiffractionGrating3f, Coherence3f, Frame3f, Transform4f, PLTBeam3f = DiffractionGrating3f, Coherence3f, Frame3f, Transform, PLTBeam3f
    ScalarVector2f, ScalarVector3f, ScalarVector2i = ScalarVector2f, ScalarVector3f, ScalarVector2i


scalar_rgb = scalar_spectral = _Namespace

_SUPPORTED_VARIANTS = ("cuda_spectral", "cuda_ad_spectral", "llvm_spectral", "llvm_ad_spectral", "scalar_spectral",
                       "cuda_spectral_polarized", "cuda_ad_spectral_polarized", "llvm_spectral_polarized",
                       "llvm_ad_spectral_polarized", "scalar_spectral_polarized",
                       "cuda_rgb", "cuda_ad_rgb", "llvm_rgb", "llvm_ad_rgb",
                       "cuda_rgb_polarized", "cuda_ad_rgb_polarized", "llvm_rgb_polarized", "llvm_ad_rgb_polarized")
_variant = "cuda_spectral"
is_spectral, is_polarized, is_rgb = True, False, False


def set_variant(*names):
    """mi.set_variant: the 4-wavelength spectral variants, unpolarized and polarized (Spectrum = Mueller matrix; what
    scripts/rendering/main-headless.py:128-130 selects); every accepted name runs the same CUDA path, the `_polarized`
    suffix switches the kernels to Mueller arithmetic (PLTB_RENDER_POLARIZED); `_rgb` selects the RGB variants (PLTB_RENDER_RGB).
    A Scene remembers the colour mode it was loaded under (as a Mitsuba scene does) and must be rendered in it."""
    global _variant, is_polarized, is_spectral, is_rgb
    for n in names:
        if n in _SUPPORTED_VARIANTS:
            _variant = n
            is_polarized = n.endswith("_polarized")
            # the `*_rgb` variants (the second mode of main-headless.py:131-133): Spectrum = Color3f, scenes are flattened with RGB
            # textures and rendered with PLTB_RENDER_RGB (roughgrating's RGB fallback, roughgrating.cpp:716-839)
            is_rgb = "_rgb" in n
            is_spectral = not is_rgb
            return
    raise ImportError(f"Requested an unsupported variant {names!r}. The following variants are available: "
                      f"{', '.join(_SUPPORTED_VARIANTS)}.")


def variant():
    return _variant


# ---------------------------------------------------------------------------------------------------------------------
# Device context (one per process / GPU)
# ---------------------------------------------------------------------------------------------------------------------
class Context:
    """Owner of a `pltb_context` (stream, film accumulator, wavefront queues, NCCL communicator)."""

    def __init__(self, device=-1, rank=0, world=1, nccl_unique_id=None):
        L = capi.lib()
        self._h = C.c_void_p()
        buf = None
        if nccl_unique_id is not None:
            buf = (C.c_char * 128).from_buffer_copy(bytes(nccl_unique_id))
        capi.check(L.pltb_context_create(device, rank, world, buf, C.byref(self._h)))
        self.rank, self.world, self.device = rank, world, device

    def close(self):
        if getattr(self, "_h", None):
            capi.lib().pltb_context_destroy(self._h)
            self._h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass


_context = None


def context():
    """Process-wide default context on the current CUDA device (rank 0 of 1 unless init_distributed() ran)."""
    global _context
    if _context is None:
        _context = Context(int(os.environ.get("LOCAL_RANK", "-1")) if "LOCAL_RANK" in os.environ else -1)
    return _context


def init_distributed():
    """One process per GPU (torchrun): build the NCCL communicator used for the single film reduce per render
    (SURVEY.md 8e).  torch.distributed is only the plumbing that ships the 128-byte ncclUniqueId to the other ranks."""
    global _context
    import torch
    import torch.distributed as dist
    rank, world = int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", str(rank)))
    if world == 1:
        _context = Context(local)
        return _context
    torch.cuda.set_device(local)
    if not dist.is_initialized():
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    uid = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        raw = (C.c_char * 128)()
        torch.zeros(1, device="cuda")  # make sure libnccl (torch's) and a CUDA context are live
        capi.check(capi.lib().pltb_nccl_unique_id(raw))
        uid = torch.frombuffer(bytearray(raw.raw), dtype=torch.uint8).clone()
    uid = uid.cuda()
    dist.broadcast(uid, 0)
    _context = Context(local, rank, world, bytes(uid.cpu().numpy().tobytes()))
    return _context


# ---------------------------------------------------------------------------------------------------------------------
# Scene
# ---------------------------------------------------------------------------------------------------------------------
class Scene:
    """Loaded scene: host-side flattened SoA (`FlatScene`) plus its lazily created device copy."""

    def __init__(self, flat: FlatScene, tree=None):
        self.flat, self.tree = flat, tree
        self._dev, self._ctx = None, None
        self._integrator = None

    # -- mitsuba.Scene surface ------------------------------------------------------------------------------------
    def integrator(self):
        if self._integrator is None and self.flat.integrator_props is not None:
            self._integrator = _make_integrator(self.flat.integrator_props)
        return self._integrator

    def sensors(self):
        return list(range(len(self.flat.sensors)))

    def film_size(self):
        d = self.flat.desc
        return d.film.width * d.n_sensors, d.film.height

    # -- device copy ---------------------------------------------------------------------------------------------------
    def device(self, ctx=None):
        ctx = ctx or context()
        if self._dev is None or self._ctx is not ctx:
            self.release()
            h = C.c_void_p()
            capi.check(capi.lib().pltb_scene_create(ctx._h, C.byref(self.flat.desc), C.byref(h)))
            self._dev, self._ctx = h, ctx
        return self._dev

    def update_sensor(self, index, to_world):
        m = np.ascontiguousarray(np.asarray(to_world, dtype=np.float64).reshape(16).astype(np.float32))
        self.flat.sensors[index].to_world[:] = list(m)
        self.flat.c_sensors[index].to_world[:] = list(m)
        if self._dev is not None:
            capi.check(capi.lib().pltb_scene_update_sensor(self._dev, index, capi.fptr(m)))

    def release(self):
        if self._dev is not None and self._ctx is not None and self._ctx._h:
            capi.lib().pltb_scene_release(self._dev)
        self._dev = None

    def __del__(self):
        try:
            self.release()
        except Exception:
            pass


class SceneParameters(dict):
    """mi.traverse(scene): the parameters the PLT scripts mutate at run time -- sensor `to_world`
    (scripts/rendering/disk_animation.py:113-123, main-headless.py:40-50)."""

    def __init__(self, scene):
        super().__init__()
        self._scene, self._dirty = scene, set()
        for i, sid in enumerate(scene.flat.sensor_ids):
            key = f"{sid or ('sensor' if i == 0 else f'sensor_{i}')}.to_world"
            m = np.array(list(scene.flat.sensors[i].to_world), dtype=np.float64).reshape(4, 4)
            dict.__setitem__(self, key, Transform(m))
            self._index = getattr(self, "_index", {})
            self._index[key] = i

    def __setitem__(self, key, value):
        if key not in self:
            raise KeyError(f'traverse(): unknown or read-only parameter "{key}"')
        dict.__setitem__(self, key, value)
        self._dirty.add(key)

    def update(self, values=None):
        if values:
            for k, v in values.items():
                self[k] = v
        for key in sorted(self._dirty):
            v = self[key]
            self._scene.update_sensor(self._index[key], v.matrix if isinstance(v, Transform) else v)
        self._dirty.clear()


def traverse(scene):
    return SceneParameters(scene)


# ---------------------------------------------------------------------------------------------------------------------
# Integrators
# ---------------------------------------------------------------------------------------------------------------------
def split_subrenders(W, H, spp):
    """The reference raises above 2^32 samples per wavefront (common.py:358-363); larger renders become n equal sub-renders
    with seeds seed, seed+1, ... whose raw accumulators are summed.  Returns (n, spp per sub-render)."""
    n_sub = 1
    while W * H * (-(-spp // n_sub)) > (1 << 32):
        n_sub += 1
    if spp % n_sub:
        raise RuntimeError(f"render(): spp={spp} exceeds the 2^32 wavefront limit and is not divisible into "
                           f"{n_sub} equal sub-renders")
    return n_sub, spp // n_sub


def shard_spp(spp, rank, world):
    """Sample-index range [lo, hi) of every pixel rendered by `rank`: contiguous, disjoint, covering."""
    return (rank * spp) // world, ((rank + 1) * spp) // world


_integrator_registry = {}


def register_integrator(name, constructor):
    """mi.register_integrator(name, lambda props: Cls(props)) (include/mitsuba/python/python.h:55-90)"""
    _integrator_registry[name] = constructor


class PLTIntegrator:
    """The `plt` plugin (scripts/rendering/integrators/plt.py:13-48): same property names and defaults; `render` is
    ADIntegrator.render (common.py:46-110) executed by the CUDA wavefront kernels."""

    def __init__(self, props=None):
        props = props if props is not None else Properties("plt")
        g = lambda k, d: props.get(k, d)
        self.max_depth = int(g("max_depth", 16))
        self.rr_depth = int(g("rr_depth", 4))
        self.sample_pass_weight = bool(g("sample_pass_weight", False))
        self.emissive_sourcing_area = float(g("emissive_sourcing_area", 1e-4))
        self.distant_sourcing_area = float(g("distant_sourcing_area", 1e-7))
        self.max_angular_spread = float(g("max_angular_spread", 1e-7))
        self.hide_emitters = bool(g("hide_emitters", False))
        # ADIntegrator.__init__ (common.py:20-44): max_depth -1 == infinity is clamped to the ABI bound
        if self.max_depth < 0 and self.max_depth != -1:
            raise RuntimeError('"max_depth" must be set to -1 (infinite) or a value >= 0')
        if self.max_depth == -1 or self.max_depth > capi.PLTB_MAX_DEPTH:
            import warnings   # the reference has no bound (dr.alloc_local(BounceData3f, max_depth), plt.py:60); the ABI does
            warnings.warn(f"plt: max_depth={self.max_depth} exceeds the bound of this build (PLTB_MAX_DEPTH = {capi.PLTB_MAX_DEPTH}); "
                          f"paths are truncated at {capi.PLTB_MAX_DEPTH} vertices, which biases the image relative to the reference",
                          RuntimeWarning, stacklevel=3)
            self.max_depth = capi.PLTB_MAX_DEPTH
        if self.rr_depth <= 0:
            raise RuntimeError('"rr_depth" must be set to a value greater than zero!')
        self._last_stats = None
        self._cancel_ctx = None

    def cancel(self):
        if self._cancel_ctx is not None:
            capi.lib().pltb_cancel(self._cancel_ctx._h)

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True, *, ctx=None, dump=None, crop=None,
               reduce=True, profile=False, download=True, out=None, _stokes=False):
        """Returns the developed image (H, W, 3) float32, or the raw RGBW accumulator (H, W, 4) if develop=False.
        `dump`: optional dict that receives per-sample 'wavelengths', 'radiance' and 'n_rays' arrays (parity mode).
        Multi-GPU: the spp range is sharded over the ranks of `ctx`; rank 0 returns the image, the others None."""
        if not isinstance(scene, Scene):
            raise TypeError("render(): expected a Scene")
        if sensor != 0:
            raise RuntimeError("render(): use a 'batch' sensor for multi-sensor scenes (sensor index must be 0)")
        ctx = ctx or context()
        flat = scene.flat
        W, H = scene.film_size()
        spp = int(spp) if spp else int(flat.spp)
        if spp <= 0:
            raise RuntimeError("render(): spp must be positive")
        dev = scene.device(ctx)
        L = capi.lib()
        # wavefront limit of the reference (common.py:358-363); larger renders are split into sub-renders with
        # consecutive seeds whose raw accumulators are summed (SURVEY.md 8b "Seed/pass convention")
        n_sub, sub_spp = split_subrenders(W, H, spp)
        if flat.rgb != is_rgb:
            raise RuntimeError(f"render(): the scene was loaded for {'an RGB' if flat.rgb else 'a spectral'} variant but the active "
                               f"variant is {_variant!r}; load it again after set_variant()")
        mode = (capi.RENDER_POLARIZED if is_polarized else 0) | (capi.RENDER_STOKES if _stokes else 0) | (capi.RENDER_RGB if flat.rgb else 0)
        nch = capi.film_channels(mode)
        shape = (H, W, nch - 1 if develop else nch)
        if out is not None:
            if out.shape != shape or out.dtype != np.float32 or not out.flags.c_contiguous:
                raise RuntimeError(f"render(): `out` must be a C-contiguous float32 array of shape {shape}")
        elif ctx.rank == 0 or not reduce:
            out = np.empty(shape, np.float32)
        total = capi.Stats()
        self._cancel_ctx = ctx
        try:
            for k in range(n_sub):
                prm = capi.RenderParams()
                prm.seed = (int(seed) + k + flat.base_seed) & 0xFFFFFFFF
                prm.spp = sub_spp
                lo, hi = shard_spp(sub_spp, ctx.rank, ctx.world) if reduce else (0, sub_spp)
                prm.spp_begin, prm.spp_count = lo, hi - lo
                prm.max_depth, prm.rr_depth = self.max_depth, self.rr_depth
                if crop:
                    prm.crop_x, prm.crop_y, prm.crop_w, prm.crop_h = crop
                last = k == n_sub - 1
                flags = mode
                if k > 0:
                    flags |= capi.RENDER_ACCUMULATE
                if last and reduce and ctx.world > 1:
                    flags |= capi.RENDER_REDUCE
                if last and develop:
                    flags |= capi.RENDER_DEVELOP
                if profile:
                    flags |= capi.RENDER_PROFILE
                prm.flags = flags
                dmp = None
                if dump is not None and last:
                    cw, ch = (crop[2], crop[3]) if crop else (W, H)
                    n = cw * ch * prm.spp_count
                    dump["wavelengths"] = np.zeros((n, 4), np.float32)
                    dump["radiance"] = np.zeros((n, 4), np.float32)
                    dump["n_rays"] = np.zeros((n, 2), np.uint32)     # closest-hit rays, shadow rays per sample
                    dmp = capi.SampleDump(capi.fptr(dump["wavelengths"]), capi.fptr(dump["radiance"]), capi.uptr(dump["n_rays"]), n)
                    if is_polarized:   # first column of the result Mueller matrix: (n, Stokes component, wavelength)
                        dump["stokes"] = np.zeros((n, 4, 4), np.float32)
                        dmp.stokes = capi.fptr(dump["stokes"])
                st = capi.Stats()
                dst = out.ctypes.data if (last and out is not None and download) else None
                capi.check(L.pltb_render(ctx._h, dev, C.byref(prm), dst, C.byref(dmp) if dmp else None, C.byref(st)))
                if n_sub == 1:
                    total = st          # the common case: no per-field accumulation on the host path of short renders
                    continue
                for f, _ in capi.Stats._fields_:
                    if f in ("ms_kernel", "launches_kernel"):
                        for j in range(16):
                            getattr(total, f)[j] += getattr(st, f)[j]
                    else:
                        setattr(total, f, getattr(total, f) + getattr(st, f))
        finally:
            self._cancel_ctx = None
        self._last_stats = total
        return out if download else None

    @property
    def last_stats(self):
        """counters and device times of the last render() (pltb_stats as a dict; built on access)"""
        return None if self._last_stats is None else self._last_stats.as_dict()

    def sample(self, *a, **kw):
        raise RuntimeError("PLTIntegrator.sample() is inlined into the CUDA wavefront kernels; use render()")


register_integrator("plt", lambda props: PLTIntegrator(props))


class StokesIntegrator:
    """The `stokes` plugin (src/integrators/stokes.cpp:66-138) around a nested `plt` integrator -- what
    scripts/rendering/utils/render.py:17-30 builds: the result gains 12 channels S0.R S0.G S0.B ... S3.B after R G B
    (`aov_names`), i.e. render() returns (H, W, 15)."""

    def __init__(self, nested):
        if not is_polarized:
            raise RuntimeError("This integrator should only be used in polarized mode!")   # stokes.cpp:73-74
        if nested is None:
            raise RuntimeError("Must specify a sub-integrator!")                           # stokes.cpp:84-85
        if not isinstance(nested, PLTIntegrator):
            raise RuntimeError("Child objects must be of type 'SamplingIntegrator'!")       # stokes.cpp:77-78
        self.nested = nested

    def aov_names(self):
        return [f"S{i}.{c}" for i in range(4) for c in "RGB"]

    def render(self, scene, sensor=0, seed=0, spp=0, develop=True, evaluate=True, **kw):
        return self.nested.render(scene, sensor=sensor, seed=seed, spp=spp, develop=develop, evaluate=evaluate, _stokes=True, **kw)

    def cancel(self):
        self.nested.cancel()

    @property
    def last_stats(self):
        return self.nested.last_stats


def _make_integrator(props, nested=None):
    t = props.plugin_name
    if t == "stokes":
        return StokesIntegrator(nested)
    if t not in _integrator_registry:
        raise RuntimeError(f'Plugin "{t}" not found (only the `plt` integrator is on the B200 hot path)')
    return _integrator_registry[t](props)


# ---------------------------------------------------------------------------------------------------------------------
# Loading / rendering entry points
# ---------------------------------------------------------------------------------------------------------------------
def load_file(path, **defaults):
    """mi.load_file(path, **defaults) (src/core/python/xml_v.cpp:68-89): `-Dkey=value` style parameter substitution."""
    if not os.path.exists(path):
        raise RuntimeError(f'"{path}": file does not exist!')
    tree = load_file_tree(path, **{k: str(v) for k, v in defaults.items()})
    return Scene(flatten(tree, rgb=is_rgb), tree)


def load_dict(d):
    """mi.load_dict (src/core/python/xml_v.cpp:91-116): a scene dictionary -> Scene; an integrator dictionary -> integrator."""
    if not isinstance(d, dict) or "type" not in d:
        raise RuntimeError("load_dict(): expected a dictionary with a 'type' entry")
    if d["type"] == "scene":
        tree = dict_to_tree(d)
        return Scene(flatten(tree, rgb=is_rgb), tree)
    if d["type"] in _integrator_registry or d["type"] == "stokes":
        p = Properties(d["type"], **{k: v for k, v in d.items() if k != "type" and not isinstance(v, dict)})
        children = [load_dict(v) for v in d.values() if isinstance(v, dict)]
        if d["type"] == "stokes" and len(children) > 1:
            raise RuntimeError("More than one sub-integrator specified!")                   # stokes.cpp:79-80
        return _make_integrator(p, children[0] if children else None)
    raise RuntimeError(f'load_dict(): top-level plugin "{d["type"]}" is not on the plt hot path')


def render(scene, params=None, sensor=0, integrator=None, seed=0, seed_grad=0, spp=0, spp_grad=0):
    """mi.render (src/python/python/util.py:378-507), forward mode only."""
    if params is not None and not isinstance(params, SceneParameters):
        raise RuntimeError("render(): `params` must be the result of traverse(scene)")
    if integrator is None:
        integrator = scene.integrator()
    if integrator is None:
        raise RuntimeError("No integrator specified! Add an integrator in the scene description or provide an "
                           "integrator directly as argument.")
    return integrator.render(scene, sensor=sensor, seed=seed, spp=spp)


class Bitmap:
    """Minimal mi.Bitmap: wraps an (H, W, C) float32 array, or reads one from an .exr / .png / .npy file; `write` supports
    .exr (linear float32), .npy and 8-bit sRGB .png (src/core/bitmap.cpp; imageio.py)."""

    def __init__(self, array):
        if isinstance(array, (str, os.PathLike)):
            path = os.fspath(array)
            ext = path.rsplit(".", 1)[-1].lower()
            if ext == "exr":
                array = imageio.read_exr(path)[0]
            elif ext == "npy":
                array = np.load(path)
            else:
                from PIL import Image
                array = np.asarray(Image.open(path), dtype=np.float32) / 255.0
        self.array = np.asarray(array, dtype=np.float32)

    def __array__(self, dtype=None, copy=None):
        return self.array if dtype is None else self.array.astype(dtype)

    def convert(self, srgb_gamma=True):
        a = np.nan_to_num(self.array, nan=0.0, posinf=1.0, neginf=0.0).clip(0, 1)
        if srgb_gamma:
            a = np.where(a <= 0.0031308, a * 12.92, 1.055 * np.power(a, 1 / 2.4) - 0.055)
        return (a * 255 + 0.5).astype(np.uint8)

    def write(self, path):
        imageio.write_bitmap(path, self.array)


class util:
    """mi.util: the two helpers the PLT scripts use (scripts/rendering/main-headless.py:60-73)."""
    write_bitmap = staticmethod(lambda path, data, write_async=False: imageio.write_bitmap(
        path, data.array if isinstance(data, Bitmap) else data, write_async))
    convert_to_bitmap = staticmethod(lambda data, uint8_srgb=True: Bitmap(data))
