"""ctypes mirror of include/pltb.h and loader for libpltb200.so (the CUDA product library).

The product path has NO CPU fallback: if the library is missing or no CUDA device is usable, `lib()` /
`Context()` raise.
"""
import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.path.join(_HERE, "libpltb200.so")
if os.environ.get("PLTB200_LIB"):   # developer switch: an experimental build of the same library (tools/gpu.sh ab:<suffix>)
    LIB_PATH = os.path.abspath(os.environ["PLTB200_LIB"])

PLTB_ABI_VERSION = 6
PLTB_MAX_DEPTH = 16

# enums (include/pltb.h)
TEX_UNIFORM, TEX_SRGB, TEX_D65, TEX_REGULAR, TEX_IRREGULAR, TEX_BITMAP, TEX_BLACKBODY, TEX_RGB = range(8)
TEXFLAG_RGB = 1
WRAP_REPEAT, WRAP_MIRROR, WRAP_CLAMP = range(3)
BSDF_NULL, BSDF_DIFFUSE, BSDF_CONDUCTOR, BSDF_DIELECTRIC, BSDF_ROUGHCONDUCTOR, BSDF_ROUGHGRATING, BSDF_PRINCIPLED = range(7)
PRINCIPLED_HAS = {"anisotropic": 1, "metallic": 2, "clearcoat": 4, "sheen": 8, "sheen_tint": 16, "flatness": 32, "spec_tint": 64,
                  "spec_trans": 128}
BSDF_FLAG_TWOSIDED = 1
MICROFACET_BECKMANN, MICROFACET_GGX = 0, 1
SHAPE_MESH, SHAPE_SPHERE, SHAPE_RECTANGLE = range(3)
SHAPE_HAS_NORMALS, SHAPE_HAS_UV, SHAPE_FLIP_NORMALS = 1, 2, 4
EMITTER_AREA, EMITTER_DIRECTIONAL, EMITTER_CONSTANT, EMITTER_ENVMAP, EMITTER_DIRECTIONALSPOT = 0, 1, 2, 3, 4
SENSOR_PERSPECTIVE, SENSOR_ORTHOGRAPHIC = 0, 1
RFILTER_BOX, RFILTER_TENT, RFILTER_GAUSSIAN = range(3)
RENDER_DEVELOP, RENDER_REDUCE, RENDER_DEVICE_OUT, RENDER_ACCUMULATE, RENDER_NO_SPLAT, RENDER_PROFILE = 1, 2, 4, 8, 16, 32
RENDER_POLARIZED, RENDER_STOKES, RENDER_RGB = 64, 128, 256


def film_channels(flags):
    """raw accumulator channels per pixel (PLTB_FILM_CHANNELS)"""
    return 16 if flags & RENDER_STOKES else 4

f32, u32, i32, u64 = C.c_float, C.c_uint32, C.c_int32, C.c_uint64
PF, PU = C.POINTER(f32), C.POINTER(u32)


class Texture(C.Structure):
    _fields_ = [("type", u32), ("flags", u32), ("offset", u32), ("count", u32), ("width", u32), ("height", u32),
                ("wrap", u32), ("pad", u32), ("p", f32 * 8), ("to_uv", f32 * 9), ("mean", f32)]


class Bsdf(C.Structure):
    _fields_ = [("type", u32), ("flags", u32), ("tex", i32 * 4), ("u", u32 * 4), ("f", f32 * 12)]


class Shape(C.Structure):
    _fields_ = [("type", u32), ("flags", u32), ("first_face", u32), ("n_faces", u32), ("bsdf", i32), ("emitter", i32),
                ("cdf_offset", u32), ("pad", u32), ("area_sum", f32), ("area_norm", f32), ("to_world", f32 * 16),
                ("to_object", f32 * 16), ("center", f32 * 3), ("radius", f32), ("frame_n", f32 * 3), ("inv_area", f32)]


class Emitter(C.Structure):
    _fields_ = [("type", u32), ("shape", i32), ("radiance", i32), ("pad", u32), ("direction", f32 * 3),
                ("bsphere_radius", f32), ("bsphere_center", f32 * 3), ("sampling_weight", f32),
                ("env_warp", u32), ("env_data", u32), ("env_res", u32 * 2), ("env_scale", f32), ("env_pad", f32 * 3),
                ("to_world", f32 * 16), ("to_local", f32 * 16)]


class Sensor(C.Structure):
    _fields_ = [("type", u32), ("srf", i32), ("sample_to_camera", f32 * 16), ("to_world", f32 * 16),
                ("near_clip", f32), ("far_clip", f32), ("pp_offset", f32 * 2)]


class Film(C.Structure):
    _fields_ = [("width", u32), ("height", u32), ("rfilter", u32), ("pad", u32), ("radius", f32), ("stddev", f32)]


class SceneDesc(C.Structure):
    _fields_ = [("abi_version", u32), ("n_vertices", u32), ("n_faces", u32), ("n_shapes", u32), ("n_bsdfs", u32),
                ("n_textures", u32), ("n_emitters", u32), ("n_sensors", u32), ("n_tex_data", u32), ("n_area_cdf", u32),
                ("positions", PF), ("normals", PF), ("texcoords", PF), ("faces", PU), ("face_shape", PU),
                ("shapes", C.POINTER(Shape)), ("bsdfs", C.POINTER(Bsdf)), ("textures", C.POINTER(Texture)),
                ("tex_data", PF), ("emitters", C.POINTER(Emitter)), ("area_cdf", PF), ("sensors", C.POINTER(Sensor)),
                ("film", Film), ("cie_srgb", PF)]


class RenderParams(C.Structure):
    _fields_ = [("seed", u32), ("spp", u32), ("spp_begin", u32), ("spp_count", u32), ("max_depth", u32),
                ("rr_depth", u32), ("flags", u32), ("crop_x", u32), ("crop_y", u32), ("crop_w", u32), ("crop_h", u32)]


KERNEL_NAMES = ["k_raygen", "k_shade<grating>", "k_shade<other>", "k_trace", "k_classify", "k_nee_light", "k_trace<shadow>",
                "k_nee_post<other>", "k_nee_post<grating>", "k_nee_eval", "k_gather", "k_splat"]   # PLTB_K_* (include/pltb.h)


class SampleDump(C.Structure):
    _fields_ = [("wavelengths", PF), ("radiance", PF), ("n_rays", PU), ("capacity", u64), ("stokes", PF)]


class Stats(C.Structure):
    _fields_ = [("samples", u64), ("closest_rays", u64), ("shadow_rays", u64), ("kernel_launches", u64),
                ("vertices", u64), ("grating_vertices", u64), ("nee_other", u64), ("nee_grating", u64), ("passes", u64), ("nee_grating_eval", u64),
                ("ms_total", f32), ("ms_sample", f32), ("ms_nee_grating", f32),
                ("ms_nee_other", f32), ("ms_splat", f32), ("ms_develop", f32), ("ms_reduce", f32), ("ms_copy", f32),
                ("ms_kernel", f32 * 16), ("launches_kernel", u32 * 16)]

    def as_dict(self):
        d = {k: getattr(self, k) for k, _ in self._fields_ if k not in ("ms_kernel", "launches_kernel")}
        d["kernels"] = {n: {"ms": self.ms_kernel[i], "launches": self.launches_kernel[i]} for i, n in enumerate(KERNEL_NAMES)}
        return d


EXPORTS = ["pltb_nccl_unique_id", "pltb_context_create", "pltb_context_destroy", "pltb_scene_create", "pltb_scene_update_sensor",
           "pltb_scene_release", "pltb_render", "pltb_cancel", "pltb_film_device_ptr", "pltb_synchronize", "pltb_host_alloc", "pltb_host_free",
           "pltb_last_error", "pltb_abi_version", "pltb_film_develop_noalloc"]
NOALLOC_MAX_FLOATS = 4 << 20   # PLTB_NOALLOC_MAX_FLOATS: the static device scratch of pltb_film_develop_noalloc
DEVELOP_ASYNC = 1

_lib = None


def lib():
    """Load libpltb200.so; raises if the CUDA extension has not been built (no fallback)."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise RuntimeError(f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; "
                               f"g.build()'` (the plt path has no CPU fallback)")
        L = C.CDLL(LIB_PATH)
        vp = C.c_void_p
        L.pltb_context_create.argtypes = [C.c_int, C.c_int, C.c_int, vp, C.POINTER(vp)]
        L.pltb_nccl_unique_id.argtypes = [vp]
        L.pltb_context_destroy.argtypes = [vp]
        L.pltb_context_destroy.restype = None
        L.pltb_scene_create.argtypes = [vp, C.POINTER(SceneDesc), C.POINTER(vp)]
        L.pltb_scene_update_sensor.argtypes = [vp, u32, PF]
        L.pltb_scene_release.argtypes = [vp]
        L.pltb_scene_release.restype = None
        L.pltb_render.argtypes = [vp, vp, C.POINTER(RenderParams), vp, C.POINTER(SampleDump), C.POINTER(Stats)]
        L.pltb_cancel.argtypes = [vp]
        L.pltb_film_device_ptr.argtypes = [vp, C.POINTER(vp), C.POINTER(u64)]
        L.pltb_synchronize.argtypes = [vp]
        L.pltb_host_alloc.argtypes = [u64, C.POINTER(vp)]
        L.pltb_host_free.argtypes = [vp]
        L.pltb_host_free.restype = None
        L.pltb_film_develop_noalloc.argtypes = [vp, u32, u32, u32, u32, vp, vp]
        L.pltb_last_error.restype = C.c_char_p
        L.pltb_abi_version.restype = C.c_int
        if L.pltb_abi_version() != PLTB_ABI_VERSION:
            raise RuntimeError("libpltb200.so ABI version mismatch")
        _lib = L
    return _lib


def check(rc):
    if rc != 0:
        raise RuntimeError(lib().pltb_last_error().decode())


def fptr(a):
    return a.ctypes.data_as(PF)


def uptr(a):
    return a.ctypes.data_as(PU)


class PinnedArray:
    """float32 numpy array over page-locked host memory (pltb_host_alloc); freed with the object."""

    def __init__(self, shape):
        self.shape = tuple(int(x) for x in shape)
        n = int(np.prod(self.shape))
        self._ptr = C.c_void_p()
        check(lib().pltb_host_alloc(4 * n, C.byref(self._ptr)))
        self.array = np.ctypeslib.as_array(C.cast(self._ptr, C.POINTER(C.c_float)), shape=(n,)).reshape(self.shape)

    def __del__(self):
        try:
            if self._ptr:
                self.array = None
                lib().pltb_host_free(self._ptr)
                self._ptr = None
        except Exception:
            pass
