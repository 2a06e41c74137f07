/*
 * pltb.h -- C ABI of libpltb200: the B200-native forward render path of the Mitsuba-PLT `plt`
 * integrator (ZenithGD/mitsuba3-plt).
 *
 * This is the drop-in boundary.  Everything above it (XML / dict scene loading,
 * plugin parameter handling, `mi.render`) is host code; everything below it is hand-written sm_100a CUDA.
 * No C++ or torch types cross this boundary: plain pointers, sizes and POD structs only.
 *
 * Reference interfaces replaced by each entry point (paths relative to the reference tree):
 *   pltb_scene_create      <- Scene ctor + accel_init  (src/render/scene.cpp:22-96,
 *                             src/render/scene_optix.inl:303 / scene_native.inl:10) and the plugin ctors that
 *                             the XML loader instantiates (src/core/xml.cpp:1368, src/core/plugin.cpp:174)
 *   pltb_scene_update_sensor <- mi.traverse(scene)["sensor.to_world"] = ..; params.update()
 *                             (src/python/python/util.py:269, src/sensors/perspective.cpp:164-172)
 *   pltb_render            <- Integrator::render (include/mitsuba/render/integrator.h:77,
 *                             src/python/python/ad/integrators/common.py:46-110) with
 *                             PLTIntegrator.sample (scripts/rendering/integrators/plt.py:492-529) inlined
 *                             PLTB_RENDER_POLARIZED | PLTB_RENDER_STOKES: StokesIntegrator::sample (src/integrators/stokes.cpp:
 *                             88-138) around it, driven by SamplingIntegrator::render (src/render/integrator.cpp:121-367,419-491)
 *   pltb_cancel            <- Integrator::cancel (src/render/integrator.cpp:91-93)
 *   pltb_last_error        <- Throw(...) -> std::runtime_error (include/mitsuba/core/logger.h:133-160)
 *
 * Conventions: all functions return 0 on success and a negative code on failure; the message of the last
 * failure on the calling thread is returned by pltb_last_error().  All pointers are caller-owned unless
 * stated.  pltb_render is blocking; one render in flight per context; contexts are independent (one host
 * thread or process per GPU).
 */
#ifndef PLTB_H
#define PLTB_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define PLTB_ABI_VERSION 6
#define PLTB_MAX_DEPTH 16      /* upper bound on the `max_depth` plugin parameter (plt.py:37) */
#define PLTB_MAX_LOBES 9       /* diffractionGratingsMaxLobes (include/mitsuba/plt/diffractiongrating.h:24) */

/* ---- error codes -------------------------------------------------------------------------------------- */
enum {
    PLTB_OK = 0,
    PLTB_ERR_INVALID = -1,     /* bad argument / malformed scene description */
    PLTB_ERR_CUDA = -2,        /* CUDA runtime failure (message carries the CUDA error string) */
    PLTB_ERR_UNSUPPORTED = -3, /* plugin type outside the hot path (SURVEY.md 2.1 "OUT OF SCOPE" rows) */
    PLTB_ERR_CANCELLED = -4,
    PLTB_ERR_NOMEM = -5
};

/* ---- flattened scene description (host SoA arrays) ----------------------------------- */

/* Spectrum / texture plugins (src/spectra/, src/textures/bitmap.cpp) */
enum {
    PLTB_TEX_UNIFORM = 0,   /* src/spectra/uniform.cpp: p[0] = value */
    PLTB_TEX_SRGB = 1,      /* src/spectra/srgb.cpp: p[0..2] = rgb2spec coefficients */
    PLTB_TEX_D65 = 2,       /* src/spectra/d65.cpp: d65 table (offset/count, range p[3],p[4]) times srgb(p[0..2])
                               when flag bit0 is set, else the bare table */
    PLTB_TEX_REGULAR = 3,   /* src/spectra/regular.cpp: values at data[offset..offset+count), range p[3],p[4] */
    PLTB_TEX_IRREGULAR = 4, /* src/spectra/irregular.cpp: wavelengths at data[offset..+count), values after them */
    PLTB_TEX_BITMAP = 5,    /* src/textures/bitmap.cpp, spectral mode: per-texel rgb2spec coefficients
                               data[offset + 3*(y*w+x) ..], bilinear, wrap mode `wrap`, uv transform to_uv */
    PLTB_TEX_BLACKBODY = 6, /* src/spectra/blackbody.cpp: p[0] = temperature, range p[3],p[4] */
    PLTB_TEX_RGB = 7        /* RGB variants (PLTB_RENDER_RGB) only -- src/spectra/srgb.cpp:56-57 with Spectrum = Color3f: p[0..2] = the
                               linear RGB value itself.  Scenes for the RGB variants hold UNIFORM, RGB and BITMAP | PLTB_TEXFLAG_RGB
                               textures only: d65 / regular / irregular / blackbody spectra are converted to RGB at load time
                               (src/spectra/d65.cpp:140-160, src/core/xml.cpp:1176-1260) */
};
enum { PLTB_TEXFLAG_D65_PRODUCT = 1,  /* PLTB_TEX_D65: multiply the table with srgb(p[0..2]) */
       PLTB_TEXFLAG_RGB = 1 };         /* PLTB_TEX_BITMAP: texels are linear RGB triples, not rgb2spec coefficients (RGB variants) */
enum { PLTB_WRAP_REPEAT = 0, PLTB_WRAP_MIRROR = 1, PLTB_WRAP_CLAMP = 2 };

typedef struct pltb_texture {
    uint32_t type;
    uint32_t flags;
    uint32_t offset, count;   /* into tex_data (floats) */
    uint32_t width, height, wrap, pad;
    float p[8];
    float to_uv[9];           /* row-major 3x3 affine uv transform (bitmap `to_uv`) */
    float mean;               /* Texture::mean(), used by eval_1 of srgb/bitmap */
} pltb_texture;

/* BSDF plugins on the path */
enum {
    PLTB_BSDF_NULL = 0,          /* shape without a BSDF */
    PLTB_BSDF_DIFFUSE = 1,       /* src/bsdfs/diffuse.cpp        tex[0] = reflectance */
    PLTB_BSDF_CONDUCTOR = 2,     /* src/bsdfs/conductor.cpp      tex[0] = eta, tex[1] = k, tex[2] = specular_reflectance */
    PLTB_BSDF_DIELECTRIC = 3,    /* src/bsdfs/dielectric.cpp     f[0] = eta, tex[2] = specular_reflectance (-1: none),
                                                                 tex[3] = specular_transmittance (-1: none) */
    PLTB_BSDF_ROUGHCONDUCTOR = 4,/* src/bsdfs/roughconductor.cpp tex[0..2] as conductor (tex[2] may be -1),
                                                                 f[1] = alpha_u, f[2] = alpha_v, u[0] = distribution,
                                                                 u[1] = sample_visible */
    PLTB_BSDF_PRINCIPLED = 6,    /* src/bsdfs/principled.cpp
                                                                 tex[0] = base_color, f[0] = eta (from `eta` or `specular`,
                                                                 principled.cpp:217-231), f[1] = roughness, f[2] = anisotropic,
                                                                 f[3] = metallic, f[4] = clearcoat, f[5] = clearcoat_gloss,
                                                                 f[6] = sheen, f[7] = sheen_tint, f[8] = flatness,
                                                                 f[9] = spec_tint, f[10] = spec_trans (the GGX refraction lobe,
                                                                 principled.cpp:437-453,604-621,816-828; not under `twosided`),
                                                                 u[0] = PLTB_PRINCIPLED_HAS_* bits,
                                                                 u[1..3] = float bits of main_specular / clearcoat /
                                                                 diffuse_reflectance _sampling_rate */
    PLTB_BSDF_ROUGHGRATING = 5   /* src/bsdfs/roughgrating.cpp   as roughconductor, plus f[3] = grating_angle,
                                                                 f[4] = inv_period_x, f[5] = inv_period_y, f[6] = height,
                                                                 f[7] = multiplier, f[8] = coherence, u[2] = lobes,
                                                                 u[3] = DiffractionGratingType bits (0 sinusoidal,
                                                                 1 rectangular, 2 linear, | 0x10 radial) */
};
enum { PLTB_BSDF_FLAG_TWOSIDED = 1 };   /* wrapped in src/bsdfs/twosided.cpp (single nested BRDF) */
enum { PLTB_MICROFACET_BECKMANN = 0, PLTB_MICROFACET_GGX = 1 };
/* get_flag() of src/bsdfs/principledhelpers.h:95-106: the property is present and not a float equal to 0 */
enum { PLTB_PRINCIPLED_HAS_ANISOTROPIC = 1, PLTB_PRINCIPLED_HAS_METALLIC = 2, PLTB_PRINCIPLED_HAS_CLEARCOAT = 4,
       PLTB_PRINCIPLED_HAS_SHEEN = 8, PLTB_PRINCIPLED_HAS_SHEEN_TINT = 16, PLTB_PRINCIPLED_HAS_FLATNESS = 32,
       PLTB_PRINCIPLED_HAS_SPEC_TINT = 64, PLTB_PRINCIPLED_HAS_SPEC_TRANS = 128 };

typedef struct pltb_bsdf {
    uint32_t type;
    uint32_t flags;
    int32_t tex[4];
    uint32_t u[4];
    float f[12];
} pltb_bsdf;

/* Shapes (src/shapes/{ply,cube,rectangle,sphere}.cpp) */
enum { PLTB_SHAPE_MESH = 0, PLTB_SHAPE_SPHERE = 1, PLTB_SHAPE_RECTANGLE = 2 };
enum { PLTB_SHAPE_HAS_NORMALS = 1, PLTB_SHAPE_HAS_UV = 2, PLTB_SHAPE_FLIP_NORMALS = 4 };

typedef struct pltb_shape {
    uint32_t type;
    uint32_t flags;
    uint32_t first_face, n_faces;   /* mesh: range in `faces` */
    int32_t bsdf;                   /* index into bsdfs */
    int32_t emitter;                /* index into emitters, -1 if not an emitter */
    uint32_t cdf_offset;            /* mesh emitters: offset of the unnormalised area CDF in area_cdf */
    uint32_t pad;
    float area_sum, area_norm;      /* DiscreteDistribution sum / normalization (1/sum) */
    float to_world[16], to_object[16]; /* analytic shapes, row-major 4x4 */
    float center[3], radius;        /* sphere */
    float frame_n[3], inv_area;     /* rectangle: world normal, 1/surface area */
} pltb_shape;

/* Emitters (src/emitters/{area,directional,constant,envmap}.cpp).  CONSTANT and ENVMAP are environment emitters: rays that leave the scene see
 * it (SurfaceInteraction::emitter(scene) of an invalid interaction, include/mitsuba/render/interaction.h:343-346); at most one
 * environment emitter per scene (src/render/scene.cpp:60-67). */
enum { PLTB_EMITTER_AREA = 0, PLTB_EMITTER_DIRECTIONAL = 1, PLTB_EMITTER_CONSTANT = 2, PLTB_EMITTER_ENVMAP = 3,
       PLTB_EMITTER_DIRECTIONALSPOT = 4 };   /* src/emitters/directionalspot.cpp (the fork's directional emitter with an angular spread):
                                                to_world = emitter frame (+z = direction of propagation), env_pad[0] = sin(spread_angle),
                                                radiance = irradiance, bsphere_* as directional */

typedef struct pltb_emitter {
    uint32_t type;
    int32_t shape;        /* area: owning shape */
    int32_t radiance;     /* texture index (area: radiance, directional: irradiance) */
    uint32_t pad;
    float direction[3];   /* directional: world-space direction of propagation */
    float bsphere_radius; /* directional / constant: scene bounding-sphere radius (constant: max(RayEpsilon, r (1 + RayEpsilon))) */
    float bsphere_center[3];
    float sampling_weight;
    /* envmap (src/emitters/envmap.cpp:110-243); `radiance` is the D65 whitepoint texture (m_d65) */
    uint32_t env_warp;    /* offset in tex_data of the Hierarchical2D<Float, 0> block over luminance * sin(theta): words
                             [n_levels, res_x, res_y, 0, {offset, width} per level] (integers as bit patterns), then the levels
                             (include/mitsuba/core/distr_2d.h:344-470; level >= 1 in 2x2-patch order, Level::index :772-776) */
    uint32_t env_data;    /* offset in tex_data of the (res_y, res_x, 4) array: rgb2spec coefficients + scale per texel (RGB variants:
                             linear r, g, b and an unused fourth float) */
    uint32_t env_res[2];  /* res_x = bitmap width + 1 (last column repeats the first), res_y = bitmap height */
    float env_scale;      /* `scale` plugin parameter */
    float env_pad[3];
    float to_world[16], to_local[16];   /* row-major 4x4 (only the linear part is used: transform_affine of directions) */
} pltb_emitter;

/* Sensors (src/sensors/{perspective,orthographic,batch}.cpp) */
enum { PLTB_SENSOR_PERSPECTIVE = 0, PLTB_SENSOR_ORTHOGRAPHIC = 1 };

typedef struct pltb_sensor {
    uint32_t type;
    int32_t srf;                  /* spectral response texture, -1: sample_rgb_spectrum */
    float sample_to_camera[16];   /* row-major 4x4 (inverse of perspective_projection, sensor.h:230-265) */
    float to_world[16];           /* row-major 4x4 camera-to-world */
    float near_clip, far_clip;
    float pp_offset[2];           /* film.size * principal_point_offset / crop_size */
} pltb_sensor;

/* Film + reconstruction filter (src/films/hdrfilm.cpp, src/rfilters/{gaussian,tent,box}.cpp) */
enum { PLTB_RFILTER_BOX = 0, PLTB_RFILTER_TENT = 1, PLTB_RFILTER_GAUSSIAN = 2 };

typedef struct pltb_film {
    uint32_t width, height;   /* of ONE sensor; a batch sensor lays its n_sensors films side by side in x */
    uint32_t rfilter;
    uint32_t pad;
    float radius;             /* filter radius */
    float stddev;             /* gaussian */
} pltb_film;

typedef struct pltb_scene_desc {
    uint32_t abi_version;
    uint32_t n_vertices, n_faces, n_shapes, n_bsdfs, n_textures, n_emitters, n_sensors;
    uint32_t n_tex_data, n_area_cdf;
    const float *positions;      /* 3 * n_vertices */
    const float *normals;        /* 3 * n_vertices (unused entries zero) */
    const float *texcoords;      /* 2 * n_vertices (unused entries zero) */
    const uint32_t *faces;       /* 3 * n_faces, global vertex indices */
    const uint32_t *face_shape;  /* n_faces, owning shape */
    const pltb_shape *shapes;
    const pltb_bsdf *bsdfs;
    const pltb_texture *textures;
    const float *tex_data;
    const pltb_emitter *emitters;
    const float *area_cdf;
    const pltb_sensor *sensors;  /* n_sensors >= 1; > 1 == `batch` sensor */
    pltb_film film;
    /* colour tables shared by the film conversion (include/mitsuba/core/spectrum.h:126-181) */
    const float *cie_srgb;       /* 6 * 95 on [360, 830] nm, 5 nm step: linear Rec.709 matching functions R, G, B (film
                                    conversion), then the CIE 1931 Y function (luminance() of the principled BSDF), then X and Z
                                    (cie1931_xyz of the roughgrating RGB fallback, roughgrating.cpp:751) */
} pltb_scene_desc;

/* ---- render call -------------------------------------------------------------------------------------------- */

enum {
    PLTB_RENDER_DEVELOP = 1,      /* write film_out as developed RGB (H*W*3), else raw RGBW accumulator (H*W*4) */
    PLTB_RENDER_REDUCE = 2,       /* sum the accumulator over all ranks of the context before develop */
    PLTB_RENDER_DEVICE_OUT = 4,   /* film_out is a device pointer */
    PLTB_RENDER_ACCUMULATE = 8,   /* do not clear the accumulator first (multi-pass renders) */
    PLTB_RENDER_NO_SPLAT = 16,    /* skip the film (per-sample dump only) */
    PLTB_RENDER_PROFILE = 32,     /* bracket every kernel with CUDA events and fill the per-kernel ms_* fields of pltb_stats */
    PLTB_RENDER_POLARIZED = 64,   /* the `*_spectral_polarized` variants: Spectrum = 4x4 Mueller matrix of 4 wavelengths
                                     (include/mitsuba/core/spectrum.h:60-95); the film still receives unpolarized_spectrum(L) */
    PLTB_RENDER_STOKES = 128,     /* with POLARIZED: the `stokes` wrapper integrator (src/integrators/stokes.cpp:88-138): 12 more film
                                     channels S0.R S0.G S0.B ... S3.B after R G B W (raw film H*W*16, developed H*W*15) */
    PLTB_RENDER_RGB = 256         /* the `*_rgb` / `*_rgb_polarized` variants: Spectrum = Color3f, no wavelength
                                     sampling (common.py:297-299), film value = L * weight without the CIE integration (:386-391),
                                     roughgrating takes its RGB fallback (roughgrating.cpp:499-506,716-839,1027-1031) at the path's
                                     three `sampling_wavelengths` (plt.py:64-70), which become live draws.  The scene must have been
                                     flattened for an RGB variant (see PLTB_TEX_RGB); combines with POLARIZED and STOKES */
};
#define PLTB_FILM_CHANNELS(flags) (((flags) & PLTB_RENDER_STOKES) ? 16 : 4)   /* raw accumulator channels per pixel */

typedef struct pltb_render_params {
    uint32_t seed;             /* mi.render(seed=) */
    uint32_t spp;              /* samples per pixel of the WHOLE render (defines lane index idx = pixel*spp + s) */
    uint32_t spp_begin;        /* this call renders s in [spp_begin, spp_begin + spp_count) */
    uint32_t spp_count;
    uint32_t max_depth;        /* plt.py:37 */
    uint32_t rr_depth;         /* plt.py:38 */
    uint32_t flags;
    /* optional pixel window for dumps / crops (all zero = whole film) */
    uint32_t crop_x, crop_y, crop_w, crop_h;
} pltb_render_params;

/* Per-sample dump for parity checks: one record per (pixel in window, s in the spp range),
 * ordered as pixel-major then sample index. */
typedef struct pltb_sample_dump {
    float *wavelengths;   /* 4 per sample, may be NULL (PLTB_RENDER_RGB: the three sampling_wavelengths of the path, third one repeated) */
    float *radiance;      /* 4 per sample (PLTIntegrator.sample result, before the sensor weight), may be NULL */
    uint32_t *n_rays;     /* 2 per sample: closest-hit rays, shadow rays; may be NULL */
    uint64_t capacity;    /* number of samples the arrays can hold */
    float *stokes;        /* PLTB_RENDER_POLARIZED only, 16 per sample, may be NULL: first column of the result Mueller matrix,
                             [component S0..S3][wavelength] (after the sensor-frame rotation when PLTB_RENDER_STOKES is set) */
} pltb_sample_dump;

/* kernels of one wavefront pass, in launch order (csrc/pltb_kernels.cu) */
enum { PLTB_K_RAYGEN = 0, PLTB_K_SHADE_GRATING, PLTB_K_SHADE_OTHER, PLTB_K_TRACE, PLTB_K_CLASSIFY, PLTB_K_NEE_LIGHT,
       PLTB_K_TRACE_SHADOW, PLTB_K_NEE_POST_OTHER, PLTB_K_NEE_POST_GRATING, PLTB_K_NEE_EVAL, PLTB_K_GATHER, PLTB_K_SPLAT,
       PLTB_K_COUNT };

typedef struct pltb_stats {
    uint64_t samples;
    uint64_t closest_rays;
    uint64_t shadow_rays;
    uint64_t kernel_launches;
    uint64_t vertices;          /* recorded bounce vertices (sum of path depths) */
    uint64_t grating_vertices;  /* of which on a roughgrating BSDF */
    uint64_t nee_other, nee_grating; /* NEE work items processed by the two k_nee launches (queue sizes) */
    uint64_t passes;            /* wavefront passes the render was split into */
    uint64_t nee_grating_eval;  /* roughgrating NEE items that reached a lobe window and had their Bessel terms evaluated */
    float ms_total;             /* device time of the whole call (CUDA events) */
    /* per-kernel device time, summed over passes; filled only with PLTB_RENDER_PROFILE */
    float ms_sample, ms_nee_grating, ms_nee_other, ms_splat, ms_develop, ms_reduce, ms_copy;
    /* per CUDA kernel (PLTB_K_*): device time summed over its launches and the number of launches; PLTB_RENDER_PROFILE only */
    float ms_kernel[16];
    uint32_t launches_kernel[16];
} pltb_stats;

typedef struct pltb_context pltb_context;
typedef struct pltb_scene pltb_scene;

/* device < 0: use the current CUDA device.  world > 1: nccl_unique_id points to the 128-byte ncclUniqueId that rank 0
 * obtained from pltb_nccl_unique_id() and shared with the other ranks (e.g. through torch.distributed). */
int pltb_nccl_unique_id(void *out128);
int pltb_context_create(int device, int rank, int world, const void *nccl_unique_id, pltb_context **out);
void pltb_context_destroy(pltb_context *ctx);

int pltb_scene_create(pltb_context *ctx, const pltb_scene_desc *desc, pltb_scene **out);
int pltb_scene_update_sensor(pltb_scene *scene, uint32_t sensor_index, const float to_world[16]);
void pltb_scene_release(pltb_scene *scene);

int pltb_render(pltb_context *ctx, const pltb_scene *scene, const pltb_render_params *params,
                float *film_out, pltb_sample_dump *dump, pltb_stats *stats);
int pltb_cancel(pltb_context *ctx);

/* Device-side helpers used by bench.py's roofline leg and by callers that keep the film on the GPU. */
int pltb_film_device_ptr(pltb_context *ctx, float **accum_rgbw, uint64_t *n_floats);
int pltb_synchronize(pltb_context *ctx);
/* Page-locked host memory for `film_out` / dump buffers: the device-to-host copy of an image then runs at PCIe speed instead of
 * through the driver's staging of pageable memory (the returned tensor of Integrator::render owns such a buffer). */
int pltb_host_alloc(uint64_t bytes, void **out);
void pltb_host_free(void *ptr);

/* Develop a raw film accumulator (width * height * channels floats, channels = 4: R G B W, or 16: + the 12 Stokes AOVs, as
 * pltb_render writes it without PLTB_RENDER_DEVELOP) into width * height * (channels - 1) floats -- every channel but W divided
 * by (W == 0 ? 1 : W), the arithmetic of pltb_render's PLTB_RENDER_DEVELOP step -- WITHOUT any device allocation.
 * No counterpart in the reference: added for callers that cannot call cudaMalloc (CUDA
 * graph capture, tight serving loops, harnesses that forbid it).  `accum` and `out` may each be host or device memory;
 * `stream` is a cudaStream_t (NULL = the default stream) on the CURRENT device.  The staged input and the developed image live
 * in a file-scope static __device__ array (csrc/pltb_noalloc.cu) that every GPU loading the library holds until exit, so
 * width * height * channels must not exceed PLTB_NOALLOC_MAX_FLOATS (PLTB_ERR_NOMEM with a message otherwise; 4 Mi floats =
 * 1024 x 1024 RGBW or 512 x 512 Stokes; 2 x 16 MiB of device memory).  All calls on a device share that array: calls that
 * overlap on the GPU (different streams, or PLTB_DEVELOP_ASYNC calls still in flight) race.  Without PLTB_DEVELOP_ASYNC the
 * stream is synchronised before returning; with it the caller synchronises before reading `out`. */
#define PLTB_NOALLOC_MAX_FLOATS (4u << 20)
enum { PLTB_DEVELOP_ASYNC = 1 };
int pltb_film_develop_noalloc(const float *accum, uint32_t width, uint32_t height, uint32_t channels, uint32_t flags,
                              float *out, void *stream);

const char *pltb_last_error(void);
int pltb_abi_version(void);

#ifdef __cplusplus
}
#endif
#endif /* PLTB_H */
