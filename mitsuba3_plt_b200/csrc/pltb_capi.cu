// pltb_capi.cu -- the C ABI of include/pltb.h: context / scene lifetime, scene upload, render orchestration,
// NCCL film reduce.  No CPU rendering path exists in this library: every entry point that computes launches CUDA kernels.
#include <cuda_runtime.h>
#include <dlfcn.h>
#include <nvtx3/nvToolsExt.h>   // header-only NVTX v3: ranges per wavefront stage for nsys / ncu --nvtx (no-ops without a tool attached)

#include <atomic>
#include <cmath>
#include <mutex>
#include <cstdio>
#include <cstdlib>
#include <cstring>
#include <string>
#include <vector>

#include "pltb_bvh.h"
#include "pltb_launch.h"

using namespace pltb;

struct NvtxRange { explicit NvtxRange(const char *name) { nvtxRangePushA(name); } ~NvtxRange() { nvtxRangePop(); } };

static thread_local std::string g_err;
static int fail(int code, const std::string &msg) { g_err = msg; return code; }
#define CU(call)                                                                                                     \
    do {                                                                                                             \
        cudaError_t e_ = (call);                                                                                     \
        if (e_ != cudaSuccess)                                                                                       \
            return fail(PLTB_ERR_CUDA, std::string(#call) + ": " + cudaGetErrorString(e_));                          \
    } while (0)

// ---- NCCL, resolved at run time from the libnccl the host process already loaded (torch's) -------------------------------
typedef struct ncclComm *ncclComm_t;
struct NcclUid { char b[128]; };   // ncclUniqueId (passed by value)
struct NcclApi {
    int (*GetUniqueId)(void *) = nullptr;
    int (*CommInitRank)(ncclComm_t *, int, NcclUid, int) = nullptr;
    int (*AllReduce)(const void *, void *, size_t, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*Reduce)(const void *, void *, size_t, int, int, int, ncclComm_t, cudaStream_t) = nullptr;
    int (*CommDestroy)(ncclComm_t) = nullptr;
    const char *(*GetErrorString)(int) = nullptr;
    bool ok = false;
};
static NcclApi g_nccl;
static bool load_nccl() {
    if (g_nccl.ok) return true;
    void *h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL | RTLD_NOLOAD);
    if (!h) h = dlopen("libnccl.so.2", RTLD_NOW | RTLD_GLOBAL);
    if (!h) h = RTLD_DEFAULT;
    *(void **) &g_nccl.GetUniqueId = dlsym(h, "ncclGetUniqueId");
    *(void **) &g_nccl.CommInitRank = dlsym(h, "ncclCommInitRank");
    *(void **) &g_nccl.AllReduce = dlsym(h, "ncclAllReduce");
    *(void **) &g_nccl.Reduce = dlsym(h, "ncclReduce");
    *(void **) &g_nccl.CommDestroy = dlsym(h, "ncclCommDestroy");
    *(void **) &g_nccl.GetErrorString = dlsym(h, "ncclGetErrorString");
    g_nccl.ok = g_nccl.GetUniqueId && g_nccl.CommInitRank && g_nccl.Reduce && g_nccl.CommDestroy;
    return g_nccl.ok;
}

struct pltb_context {
    int device = 0, rank = 0, world = 1;
    cudaStream_t stream = nullptr;
    ncclComm_t comm = nullptr;
    std::atomic<int> cancel{ 0 };
    // film
    float *accum = nullptr; uint64_t accum_floats = 0;
    float *rgb = nullptr; uint64_t rgb_floats = 0;
    // pass state
    PassState st{};
    uint32_t st_chunk = 0, st_depth = 0;
    bool st_polarized = false;
    uint32_t st_class_mask = 0;   // classes whose queues are allocated
    std::vector<void *> st_allocs;
    unsigned long long *counters = nullptr;
    unsigned long long *h_counters = nullptr;   // page-locked: read back on the render stream, before its one synchronisation
    float *dump_wl = nullptr, *dump_L = nullptr, *dump_stokes = nullptr; uint32_t *dump_rays = nullptr; uint64_t dump_cap = 0;
    uint32_t *n_shadow = nullptr; uint32_t n_shadow_cap = 0;   // per-sample shadow-ray counts (only for pltb_sample_dump.n_rays)
    int sm_count = 148;
    cudaEvent_t ev[2] = { nullptr, nullptr };
    std::vector<cudaEvent_t> prof;   // PLTB_RENDER_PROFILE: 5 events per pass + 4 for the tail
    std::vector<cudaEvent_t> kprof;  // PLTB_RENDER_PROFILE: one event after every kernel launch (+ one at the start of a pass)
    std::vector<int> kprof_id;       // kernel id (PLTB_K_*) that ended at kprof[i]; -1 = start marker
    size_t kprof_n = 0;
};

struct pltb_scene {
    pltb_context *ctx = nullptr;
    DeviceScene dev{};
    std::vector<void *> allocs;
    std::vector<pltb_sensor> sensors;
    pltb_sensor *d_sensors = nullptr;
    uint32_t film_w = 0, film_h = 0;
    uint32_t class_mask = 0;   // material classes (BSDF plugin types) present: bit c <=> some shape uses a BSDF of type c
    bool has_rgb_tex = false, has_spectral_tex = false;   // texture kinds present: a scene is flattened for ONE colour mode
};

template <class T> static int upload(std::vector<void *> &allocs, const T *host, size_t n, const T **dev) {
    void *p = nullptr;
    size_t bytes = std::max<size_t>(n, 1) * sizeof(T);
    CU(cudaMalloc(&p, bytes));
    allocs.push_back(p);
    if (n) CU(cudaMemcpy(p, host, n * sizeof(T), cudaMemcpyHostToDevice));
    *dev = (const T *) p;
    return 0;
}

namespace pltb {
int set_last_error(int code, const char *msg) { return fail(code, msg); }   // for pltb_noalloc.cu
}

extern "C" {

const char *pltb_last_error(void) { return g_err.c_str(); }
int pltb_abi_version(void) { return PLTB_ABI_VERSION; }

int pltb_nccl_unique_id(void *out128) {
    if (!load_nccl()) return fail(PLTB_ERR_UNSUPPORTED, "NCCL is not loaded in this process (import torch first)");
    int r = g_nccl.GetUniqueId(out128);
    return r == 0 ? 0 : fail(PLTB_ERR_CUDA, "ncclGetUniqueId failed");
}

int pltb_context_create(int device, int rank, int world, const void *nccl_unique_id, pltb_context **out) {
    if (!out || world < 1 || rank < 0 || rank >= world) return fail(PLTB_ERR_INVALID, "pltb_context_create: bad arguments");
    int n = 0;
    cudaError_t e = cudaGetDeviceCount(&n);
    if (e != cudaSuccess || n == 0)
        return fail(PLTB_ERR_CUDA, std::string("no CUDA device available (the plt path has no CPU fallback): ") + cudaGetErrorString(e));
    if (device < 0) CU(cudaGetDevice(&device));
    CU(cudaSetDevice(device));
    pltb_context *c = new pltb_context();
    c->device = device; c->rank = rank; c->world = world;
    cudaDeviceProp prop;
    CU(cudaGetDeviceProperties(&prop, device));
    c->sm_count = prop.multiProcessorCount;
    CU(cudaStreamCreateWithFlags(&c->stream, cudaStreamNonBlocking));
    CU(cudaEventCreate(&c->ev[0])); CU(cudaEventCreate(&c->ev[1]));
    CU(cudaMalloc(&c->counters, 8 * sizeof(unsigned long long)));
    CU(cudaHostAlloc((void **) &c->h_counters, 8 * sizeof(unsigned long long), cudaHostAllocDefault));
    if (world > 1) {
        if (!nccl_unique_id) { delete c; return fail(PLTB_ERR_INVALID, "world > 1 needs an NCCL unique id"); }
        if (!load_nccl()) { delete c; return fail(PLTB_ERR_UNSUPPORTED, "NCCL is not loaded in this process"); }
        NcclUid uid; memcpy(uid.b, nccl_unique_id, 128);
        int r = g_nccl.CommInitRank(&c->comm, world, uid, rank);
        if (r != 0) { delete c; return fail(PLTB_ERR_CUDA, std::string("ncclCommInitRank: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error")); }
    }
    *out = c;
    return 0;
}

static void free_state(pltb_context *c) {
    for (void *p : c->st_allocs) cudaFree(p);
    c->st_allocs.clear(); c->st_chunk = 0; c->st_depth = 0; c->st_polarized = false; c->st_class_mask = 0;
}

void pltb_context_destroy(pltb_context *c) {
    if (!c) return;
    cudaSetDevice(c->device);
    cudaStreamSynchronize(c->stream);
    if (c->comm && g_nccl.CommDestroy) g_nccl.CommDestroy(c->comm);
    free_state(c);
    cudaFree(c->accum); cudaFree(c->rgb); cudaFree(c->counters); cudaFreeHost(c->h_counters); cudaFree(c->dump_wl); cudaFree(c->dump_L);
    cudaFree(c->dump_rays); cudaFree(c->n_shadow); cudaFree(c->dump_stokes);
    cudaEventDestroy(c->ev[0]); cudaEventDestroy(c->ev[1]);
    for (cudaEvent_t e : c->prof) cudaEventDestroy(e);
    for (cudaEvent_t e : c->kprof) cudaEventDestroy(e);
    cudaStreamDestroy(c->stream);
    delete c;
}

int pltb_scene_create(pltb_context *ctx, const pltb_scene_desc *d, pltb_scene **out) {
    if (!ctx || !d || !out) return fail(PLTB_ERR_INVALID, "pltb_scene_create: null argument");
    if (d->abi_version != PLTB_ABI_VERSION) return fail(PLTB_ERR_INVALID, "pltb_scene_create: ABI version mismatch");
    if (d->n_sensors < 1) return fail(PLTB_ERR_INVALID, "pltb_scene_create: scene has no sensor");
    for (uint32_t i = 0; i < d->n_bsdfs; ++i) {
        if (d->bsdfs[i].type > PLTB_BSDF_PRINCIPLED) return fail(PLTB_ERR_UNSUPPORTED, "unsupported BSDF type");
        if (d->bsdfs[i].type == PLTB_BSDF_ROUGHGRATING && (d->bsdfs[i].u[2] < 1 || d->bsdfs[i].u[2] > PLTB_MAX_LOBES))
            return fail(PLTB_ERR_INVALID, "roughgrating: lobes must be in [1, 9]");
    }
    for (uint32_t i = 0; i < d->n_faces * 3; ++i)
        if (d->faces[i] >= d->n_vertices) return fail(PLTB_ERR_INVALID, "face index out of range");
    bool has_rgb_tex = false, has_spectral_tex = false;
    for (uint32_t i = 0; i < d->n_textures; ++i) {
        const pltb_texture &t = d->textures[i];
        if (t.type > PLTB_TEX_RGB) return fail(PLTB_ERR_UNSUPPORTED, "unsupported texture type");
        if (t.type == PLTB_TEX_RGB || (t.type == PLTB_TEX_BITMAP && (t.flags & PLTB_TEXFLAG_RGB))) has_rgb_tex = true;
        else if (t.type != PLTB_TEX_UNIFORM) has_spectral_tex = true;
    }
    if (has_rgb_tex && has_spectral_tex)
        return fail(PLTB_ERR_INVALID, "pltb_scene_create: RGB (PLTB_TEX_RGB / RGB bitmap) and spectral textures in one scene");
    for (uint32_t i = 0; i < d->n_emitters; ++i) {
        const pltb_emitter &e = d->emitters[i];
        if (e.type > PLTB_EMITTER_DIRECTIONALSPOT) return fail(PLTB_ERR_UNSUPPORTED, "unsupported emitter type");
        if (e.type == PLTB_EMITTER_ENVMAP &&
            (e.env_res[0] < 3 || e.env_res[1] < 3 || (uint64_t) e.env_data + 4ull * e.env_res[0] * e.env_res[1] > d->n_tex_data ||
             e.env_warp >= d->n_tex_data))
            return fail(PLTB_ERR_INVALID, "envmap: data / warp blocks outside tex_data");
    }
    // leaf references hold the first primitive in 24 bits; the traversal stack has kBvhStack entries
    if ((uint64_t) d->n_faces + d->n_shapes >= (1u << 24))
        return fail(PLTB_ERR_UNSUPPORTED, "scenes with 2^24 or more primitives are not supported by the BVH leaf encoding");
    HostBVH bvh;
    build_bvh(d, bvh);
    if (bvh.stack_need > (uint32_t) kBvhStack)
        return fail(PLTB_ERR_UNSUPPORTED, "BVH too deep for the traversal stack (" + std::to_string(bvh.stack_need) + " > " +
                                          std::to_string(kBvhStack) + " entries)");
    CU(cudaSetDevice(ctx->device));
    pltb_scene *s = new pltb_scene();
    s->ctx = ctx;
    s->has_rgb_tex = has_rgb_tex; s->has_spectral_tex = has_spectral_tex;
    DeviceScene &ds = s->dev;
    int rc = 0;
    const float *nodes = nullptr, *tris = nullptr;
    rc |= upload(s->allocs, bvh.nodes.data(), bvh.nodes.size(), &nodes);
    rc |= upload(s->allocs, bvh.tri_data.data(), bvh.tri_data.size(), &tris);
    ds.bvh_nodes = (const float4 *) nodes; ds.tri_data = (const float4 *) tris;
    ds.n_nodes = bvh.n_nodes; ds.n_prims = bvh.n_prims;
    rc |= upload(s->allocs, d->positions, 3 * (size_t) d->n_vertices, &ds.positions);
    rc |= upload(s->allocs, d->normals, 3 * (size_t) d->n_vertices, &ds.normals);
    rc |= upload(s->allocs, d->texcoords, 2 * (size_t) d->n_vertices, &ds.texcoords);
    rc |= upload(s->allocs, d->faces, 3 * (size_t) d->n_faces, &ds.faces);
    rc |= upload(s->allocs, d->face_shape, (size_t) d->n_faces, &ds.face_shape);
    rc |= upload(s->allocs, d->shapes, d->n_shapes, &ds.shapes);
    rc |= upload(s->allocs, d->bsdfs, d->n_bsdfs, &ds.bsdfs);
    rc |= upload(s->allocs, d->textures, d->n_textures, &ds.textures);
    rc |= upload(s->allocs, d->tex_data, d->n_tex_data, &ds.tex_data);
    rc |= upload(s->allocs, d->emitters, d->n_emitters, &ds.emitters);
    rc |= upload(s->allocs, d->area_cdf, d->n_area_cdf, &ds.area_cdf);
    rc |= upload(s->allocs, d->sensors, d->n_sensors, &ds.sensors);
    rc |= upload(s->allocs, d->cie_srgb, 6 * 95, &ds.cie_srgb);
    if (rc) { pltb_scene_release(s); return PLTB_ERR_CUDA; }
    s->d_sensors = (pltb_sensor *) ds.sensors;
    s->sensors.assign(d->sensors, d->sensors + d->n_sensors);
    ds.film = d->film;
    ds.n_faces = d->n_faces; ds.n_shapes = d->n_shapes; ds.n_emitters = d->n_emitters; ds.n_sensors = d->n_sensors;
    ds.emitter_pmf = d->n_emitters ? 1.f / (float) d->n_emitters : 0.f;
    ds.env = -1;
    for (uint32_t i = 0; i < d->n_emitters; ++i)
        if (d->emitters[i].type == PLTB_EMITTER_CONSTANT || d->emitters[i].type == PLTB_EMITTER_ENVMAP) ds.env = (int32_t) i;
    // Simpson nodes of math::bessel_j_small (include/mitsuba/core/math.h:204-223, 316-322), float ops as the reference
    {
        const float a = 0.f, b = kPi, h = (b - a) / (float) (20 - 1);
        for (int i = 0; i < 20; ++i) { float t = a + h * (float) i; ds.bessel.t[i] = t; ds.bessel.sin_t[i] = sinf(t); }
        ds.bessel.t[0] = a; ds.bessel.sin_t[0] = sinf(a); ds.bessel.t[19] = b; ds.bessel.sin_t[19] = sinf(b);
        ds.bessel.h_over_3 = h / 3.0f;
        // the same rule tabulated as cubic Hermite segments (BesselTables::tab): values and derivatives in double from the float nodes
        static std::vector<float4> tab;   // scene-independent; built once per process
        static std::mutex tab_mutex;
        {
            std::lock_guard<std::mutex> lock(tab_mutex);
            if (tab.empty()) {
                tab.resize((size_t) kBesselTabOrders * kBesselTabN);
                const double hs = 1.0 / kBesselTabInvH, pi = 3.14159265358979323846;
                auto eval = [&](int l, double x, double &f, double &df) {
                    double sum = 0, dsum = 0;
                    for (int i = 0; i < 20; ++i) {
                        const double w = (i == 0 || i == 19) ? 1.0 : ((i & 1) ? 4.0 : 2.0), ti = ds.bessel.t[i], si = ds.bessel.sin_t[i];
                        const double arg = (double) l * ti - x * si;
                        sum += w * cos(arg); dsum += w * si * sin(arg);
                    }
                    const double c = (double) kInvPi * (double) ds.bessel.h_over_3;
                    const double sm = c * sum, dsm = c * dsum;
                    const double wf = 1.0 / (1.0 + exp(-1.5 * (fabs(x) - (45.0 + l))));
                    f = (1.0 - wf) * sm; df = (1.0 - wf) * dsm;
                    if (wf > 1e-30 && x > 1e-3) {   // blend with the asymptotic form (math.h:316-347)
                        const double amp = sqrt(2.0 / (pi * x)), ph = x - pi * l / 2.0 - pi / 4.0, as = amp * cos(ph);
                        const double das = -amp * sin(ph) - 0.5 / x * as, dwf = 1.5 * wf * (1.0 - wf);
                        f += wf * as; df += wf * das + dwf * (as - sm);
                    }
                };
                for (int l = 1; l <= kBesselTabOrders; ++l)
                    for (int i = 0; i < kBesselTabN; ++i) {
                        double f0, d0, f1, d1;
                        eval(l, i * hs, f0, d0); eval(l, (i + 1) * hs, f1, d1);
                        tab[(size_t) (l - 1) * kBesselTabN + i] = make_float4((float) f0, (float) (hs * d0), (float) f1, (float) (hs * d1));
                    }
            }
        }
        rc |= upload(s->allocs, tab.data(), tab.size(), &ds.bessel.tab);
        if (rc) { pltb_scene_release(s); return PLTB_ERR_CUDA; }
    }
    s->film_w = d->film.width * d->n_sensors; s->film_h = d->film.height;
    for (uint32_t i = 0; i < d->n_shapes; ++i)
        s->class_mask |= 1u << material_class(d->shapes[i].bsdf >= 0 ? d->bsdfs[d->shapes[i].bsdf].type : 0u);
    *out = s;
    return 0;
}

int pltb_scene_update_sensor(pltb_scene *s, uint32_t sensor_index, const float to_world[16]) {
    if (!s || sensor_index >= s->sensors.size()) return fail(PLTB_ERR_INVALID, "pltb_scene_update_sensor: bad sensor index");
    CU(cudaSetDevice(s->ctx->device));
    memcpy(s->sensors[sensor_index].to_world, to_world, 16 * sizeof(float));
    // stream-ordered: the copy is queued behind the renders already submitted and ahead of the next one; the source is pageable
    // host memory, which the runtime stages before cudaMemcpyAsync returns, so no synchronisation is needed here
    CU(cudaMemcpyAsync(s->d_sensors + sensor_index, &s->sensors[sensor_index], sizeof(pltb_sensor), cudaMemcpyHostToDevice,
                       s->ctx->stream));
    return 0;
}

void pltb_scene_release(pltb_scene *s) {
    if (!s) return;
    cudaSetDevice(s->ctx->device);
    for (void *p : s->allocs) cudaFree(p);
    delete s;
}

int pltb_cancel(pltb_context *c) { if (c) c->cancel.store(1); return 0; }
int pltb_host_alloc(uint64_t bytes, void **out) {
    if (!out || !bytes) return fail(PLTB_ERR_INVALID, "pltb_host_alloc: bad argument");
    CU(cudaHostAlloc(out, bytes, cudaHostAllocDefault));
    return 0;
}
void pltb_host_free(void *ptr) { if (ptr) cudaFreeHost(ptr); }
int pltb_synchronize(pltb_context *c) { if (!c) return fail(PLTB_ERR_INVALID, "null context"); CU(cudaStreamSynchronize(c->stream)); return 0; }
int pltb_film_device_ptr(pltb_context *c, float **accum, uint64_t *n) {
    if (!c) return fail(PLTB_ERR_INVALID, "null context");
    if (accum) *accum = c->accum;
    if (n) *n = c->accum_floats;
    return 0;
}

static int ensure_state(pltb_context *c, uint32_t chunk, uint32_t depth, bool polarized, uint32_t class_mask) {
    if (c->st_chunk >= chunk && c->st_depth >= depth && (c->st_polarized || !polarized) && (class_mask & ~c->st_class_mask) == 0) return 0;
    class_mask |= c->st_class_mask;
    chunk = std::max(chunk, c->st_chunk); depth = std::max(depth, c->st_depth); polarized = polarized || c->st_polarized;
    free_state(c);
    auto alloc = [&](void **p, size_t bytes) -> int {
        CU(cudaMalloc(p, bytes));
        c->st_allocs.push_back(*p);
        return 0;
    };
    PassState &st = c->st;
    size_t nv = (size_t) chunk * depth;
    int rc = 0;
    rc |= alloc((void **) &st.wl, chunk * sizeof(float4));
    rc |= alloc((void **) &st.sensor_w, chunk * sizeof(float4));
    rc |= alloc((void **) &st.film_pos, chunk * sizeof(float2));
    {   // interleaved per-path streams (Strided, pltb_integrator.cuh)
        uint64_t *rec = nullptr; float4 *ray = nullptr, *thr = nullptr;
        rc |= alloc((void **) &rec, 4 * (size_t) chunk * sizeof(uint64_t));
        rc |= alloc((void **) &ray, 2 * (size_t) chunk * sizeof(float4));
        rc |= alloc((void **) &thr, 2 * (size_t) chunk * sizeof(float4));
        st.rng_state.base = rec; st.rng_inc.base = rec + 1; st.p_flags.base = rec + 2;
        st.p_ray_o.base = ray; st.p_ray_d.base = ray + 1;
        st.p_alpha.base = thr; st.p_prev.base = thr + 1;
    }
    rc |= alloc((void **) &st.n_vertices, chunk * sizeof(uint32_t));
    rc |= alloc((void **) &st.p_hit, chunk * sizeof(float4));
    rc |= alloc((void **) &st.v_mask, chunk * sizeof(uint32_t));
    rc |= alloc((void **) &st.v_w, nv * sizeof(float4));
    rc |= alloc((void **) &st.live, chunk * sizeof(uint32_t));
    rc |= alloc((void **) &st.hit_cls, chunk * sizeof(uint8_t));
    for (int k = 0; k < kClasses; ++k) {
        st.shade_q[k] = nullptr;
        if (!(class_mask >> k & 1u)) continue;
        rc |= alloc((void **) &st.shade_q[k], chunk * sizeof(uint32_t));
    }
    rc |= alloc((void **) &st.v_p_u, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_n_v, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_s_wix, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_shn_wiy, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_alpha, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_emw, nv * sizeof(float4));
    rc |= alloc((void **) &st.rec_entry, nv * sizeof(uint32_t));
    rc |= alloc((void **) &st.v_em, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_nee, nv * sizeof(float4));
    rc |= alloc((void **) &st.v_wiz_shape, nv * sizeof(float2));
    rc |= alloc((void **) &st.queue_eval, nv * sizeof(uint32_t));
    rc |= alloc((void **) &st.shadow_q, nv * sizeof(uint32_t));
    rc |= alloc((void **) &st.counts, kCountsTotal * sizeof(uint32_t));
    st.pp_alpha = st.pv_w = st.pv_alpha = st.pv_em = nullptr; st.stokes_theta = nullptr;
    if (polarized) {
        rc |= alloc((void **) &st.pp_alpha, 4 * (size_t) chunk * sizeof(float4));
        rc |= alloc((void **) &st.pv_w, 16 * nv * sizeof(float4));
        rc |= alloc((void **) &st.pv_alpha, 4 * nv * sizeof(float4));
        rc |= alloc((void **) &st.pv_em, 4 * nv * sizeof(float4));
        rc |= alloc((void **) &st.stokes_theta, (size_t) chunk * sizeof(float));
    }
    st.nv = nv;
    if (rc) { free_state(c); return rc; }
    c->st_polarized = polarized; c->st_class_mask = class_mask;
    st.chunk = chunk;
    st.counters = c->counters;
    c->st_chunk = chunk; c->st_depth = depth;
    return 0;
}

int pltb_render(pltb_context *c, const pltb_scene *scene, const pltb_render_params *prm, float *film_out, pltb_sample_dump *dump,
                pltb_stats *stats) {
    if (!c || !scene || !prm) return fail(PLTB_ERR_INVALID, "pltb_render: null argument");
    if (scene->ctx != c) return fail(PLTB_ERR_INVALID, "pltb_render: scene belongs to another context");
    if (prm->max_depth < 1 || prm->max_depth > PLTB_MAX_DEPTH) return fail(PLTB_ERR_INVALID, "max_depth must be in [1, 16]");
    if (prm->rr_depth < 1) return fail(PLTB_ERR_INVALID, "\"rr_depth\" must be set to a value greater than zero!");
    if (prm->spp == 0 || prm->spp_begin + prm->spp_count > prm->spp) return fail(PLTB_ERR_INVALID, "bad spp window");
    const uint32_t W = scene->film_w, H = scene->film_h;
    if ((uint64_t) W * H * prm->spp > (1ull << 32))
        return fail(PLTB_ERR_INVALID, "The total number of Monte Carlo samples required by this rendering task exceeds 2^32 = "
                                      "4294967296. Please use fewer samples per pixel or render using multiple passes.");
    CU(cudaSetDevice(c->device));
    c->cancel.store(0);
    cudaStream_t s = c->stream;
    const uint32_t cw = prm->crop_w ? prm->crop_w : W, ch = prm->crop_h ? prm->crop_h : H;
    if (prm->crop_x + cw > W || prm->crop_y + ch > H) return fail(PLTB_ERR_INVALID, "crop window outside the film");
    const uint64_t total = (uint64_t) cw * ch * prm->spp_count;
    const bool polarized = (prm->flags & PLTB_RENDER_POLARIZED) != 0, stokes = polarized && (prm->flags & PLTB_RENDER_STOKES) != 0;
    if ((prm->flags & PLTB_RENDER_STOKES) && !polarized)
        return fail(PLTB_ERR_INVALID, "This integrator should only be used in polarized mode!");   // stokes.cpp:73-74
    // colour mode: the scene copy passed to the kernels of this call carries it (DeviceScene::rgb)
    const bool rgb = (prm->flags & PLTB_RENDER_RGB) != 0;
    if (rgb && scene->has_spectral_tex)
        return fail(PLTB_ERR_INVALID, "PLTB_RENDER_RGB: the scene holds spectral textures (it was flattened for a spectral variant)");
    if (!rgb && scene->has_rgb_tex)
        return fail(PLTB_ERR_INVALID, "the scene holds RGB textures (it was flattened for an RGB variant): render it with PLTB_RENDER_RGB");
    if (rgb)
        for (const pltb_sensor &sn : scene->sensors)
            if (sn.srf >= 0) return fail(PLTB_ERR_INVALID, "PLTB_RENDER_RGB: sensor response functions need a spectral variant");
    DeviceScene dev_scene = scene->dev;
    dev_scene.rgb = rgb ? 1u : 0u;
    const int nch = PLTB_FILM_CHANNELS(stokes ? PLTB_RENDER_STOKES : 0);
    const uint64_t n_acc = (uint64_t) W * H * nch;
    if (c->accum_floats != n_acc) {
        cudaFree(c->accum); c->accum = nullptr;
        CU(cudaMalloc(&c->accum, n_acc * sizeof(float)));
        CU(cudaMemsetAsync(c->accum, 0, n_acc * sizeof(float), s));
        c->accum_floats = n_acc;
    }
    // samples per wavefront pass: large enough that the nearly empty late-bounce launches are amortised (their cost is the
    // latency of one thread's work, not throughput).  Measured on cbox-grating, 1024 spp: 2^22 1356, 2^23 1535, 2^24 1622,
    // 2^25 1666 Msamples/s.  State is 156 B per sample + 152 B per vertex slot (1.2 KB at max_depth 7): 2^25 samples = 41 GB of
    // the B200's 180 GB; the pass shrinks when that would exceed 30 % of the device memory.
    uint32_t chunk_log2 = 25;
    const bool state_fits = c->st_chunk && c->st_depth >= prm->max_depth && (total <= c->st_chunk || c->st_chunk >= (1u << 25)) &&
                            (c->st_polarized || !polarized);
    if (polarized && !c->st_polarized && c->st_chunk) free_state(c);   // the polarized arrays need a (smaller) pass of their own
    if (!state_fits) {   // (cudaMemGetInfo is a driver round trip: only when the state has to be (re)allocated)
        size_t free_b = 0, total_b = 0;
        if (cudaMemGetInfo(&free_b, &total_b) == cudaSuccess && total_b) {
            // polarized: + 64 B per sample (throughput column) + 384 B per vertex slot (16-float4 weight, two 4-float4 columns)
            const uint64_t per_sample = 164ull + 172ull * prm->max_depth + (polarized ? 68ull + 384ull * prm->max_depth : 0ull);
            while (chunk_log2 > 20 && (per_sample << chunk_log2) > (uint64_t) (0.3 * (double) total_b)) --chunk_log2;
        }
    }
    if (const char *e = getenv("PLTB_CHUNK_LOG2")) chunk_log2 = (uint32_t) std::min(26, std::max(10, atoi(e)));
    uint32_t chunk = (uint32_t) std::min<uint64_t>(std::max<uint64_t>(total, 1), 1ull << chunk_log2);
    if (int rc = ensure_state(c, chunk, prm->max_depth, polarized, scene->class_mask)) return rc;
    PassState st = c->st;
    if (!polarized) { st.pp_alpha = st.pv_w = st.pv_alpha = st.pv_em = nullptr; st.stokes_theta = nullptr; }
    else if (!stokes) st.stokes_theta = nullptr;
    st.nv = (size_t) c->st_chunk * c->st_depth;
    for (int k = 0; k < 4; ++k) st.L_out[k] = polarized ? st.pv_em + (size_t) k * st.nv : (k == 0 ? st.v_em : nullptr);
    st.chunk = c->st_chunk; st.max_depth = prm->max_depth; st.rr_depth = prm->rr_depth;
    float *d_wl = nullptr, *d_L = nullptr, *d_stokes = nullptr; uint32_t *d_rays = nullptr; uint64_t dcap = 0;
    st.n_shadow = nullptr;
    if (dump && (dump->wavelengths || dump->radiance || dump->n_rays || (polarized && dump->stokes))) {
        dcap = std::min<uint64_t>(dump->capacity, total);
        if (c->dump_cap < dcap) {
            cudaFree(c->dump_wl); cudaFree(c->dump_L); cudaFree(c->dump_rays); cudaFree(c->dump_stokes);
            c->dump_wl = c->dump_L = c->dump_stokes = nullptr; c->dump_rays = nullptr; c->dump_cap = 0;
            CU(cudaMalloc(&c->dump_wl, dcap * 4 * sizeof(float)));
            CU(cudaMalloc(&c->dump_L, dcap * 4 * sizeof(float)));
            CU(cudaMalloc(&c->dump_rays, dcap * 2 * sizeof(uint32_t)));
            CU(cudaMalloc(&c->dump_stokes, dcap * 16 * sizeof(float)));
            c->dump_cap = dcap;
        }
        d_wl = c->dump_wl; d_L = c->dump_L;
        if (polarized && dump->stokes) d_stokes = c->dump_stokes;
        if (dump->n_rays) {
            if (c->n_shadow_cap < st.chunk) {
                cudaFree(c->n_shadow); c->n_shadow = nullptr; c->n_shadow_cap = 0;
                CU(cudaMalloc(&c->n_shadow, (size_t) st.chunk * sizeof(uint32_t)));
                c->n_shadow_cap = st.chunk;
            }
            st.n_shadow = c->n_shadow; d_rays = c->dump_rays;
        }
    }
    uint64_t launches = 0;
    const bool profile = (prm->flags & PLTB_RENDER_PROFILE) != 0;
    const uint64_t n_pass = (total + st.chunk - 1) / st.chunk;
    if (profile) {
        while (c->prof.size() < 5 * n_pass + 4) { cudaEvent_t e; CU(cudaEventCreate(&e)); c->prof.push_back(e); }
    }
    // per-kernel events: the launchers call back after every kernel (pltb_launch.h)
    struct KHook {
        static void mark(int kernel, void *user) {
            pltb_context *c = (pltb_context *) user;
            if (c->kprof_n == c->kprof.size()) { cudaEvent_t e; if (cudaEventCreate(&e) != cudaSuccess) return; c->kprof.push_back(e); }
            cudaEventRecord(c->kprof[c->kprof_n++], c->stream);
            c->kprof_id.push_back(kernel);
        }
    };
    c->kprof_n = 0; c->kprof_id.clear();
    g_launch_hook = profile ? LaunchHook{ &KHook::mark, c } : LaunchHook{};
    struct HookReset { ~HookReset() { g_launch_hook = LaunchHook{}; } } hook_reset;
    size_t pe = 0;
#define PROF() do { if (profile) CU(cudaEventRecord(c->prof[pe++], s)); } while (0)
    CU(cudaEventRecord(c->ev[0], s));
    CU(cudaMemsetAsync(c->counters, 0, 8 * sizeof(unsigned long long), s));
    if (!(prm->flags & PLTB_RENDER_ACCUMULATE)) CU(cudaMemsetAsync(c->accum, 0, n_acc * sizeof(float), s));
    PassParams pp{};
    pp.seed = prm->seed; pp.spp = prm->spp; pp.spp_begin = prm->spp_begin; pp.spp_count = prm->spp_count;
    pp.crop_x = prm->crop_x; pp.crop_y = prm->crop_y; pp.crop_w = cw; pp.crop_h = ch;
    pp.film_w = W; pp.film_h = H; pp.n_pix = cw * ch;
    // grid of the grid-stride / persistent-warp kernels: up to grid_mult blocks of 128 threads per SM so that the tail of a launch
    // is short (cbox-grating 1024 spp: x4 1531, x8 1660, x16 1691, x32 1712, x64 1728, x128 1713 Msamples/s), but no more blocks than a pass of this
    // size can feed -- short renders (config 5: 4.8e5 samples per frame) pay for every block they launch
    static const int grid_mult = getenv("PLTB_GRID_MULT") ? std::max(1, atoi(getenv("PLTB_GRID_MULT"))) : 64;
    const uint64_t pass_samples = std::min<uint64_t>(total, st.chunk);
    const int wf_grid = (int) std::min<uint64_t>((uint64_t) c->sm_count * grid_mult,
                                                 std::max<uint64_t>(c->sm_count, (pass_samples + 511) / 512));
    const int nee_grid = wf_grid;
    const int do_splat = (prm->flags & PLTB_RENDER_NO_SPLAT) ? 0 : 1;
    const int splat_nr = dev_scene.film.rfilter == PLTB_RFILTER_BOX ? 0 : (int) ceilf(dev_scene.film.radius - .5f);
    const bool strip_splat = (splat_nr == 1 || splat_nr == 2) && !getenv("PLTB_SPLAT_DIRECT");   // as launch_splat decides
    for (uint64_t first = 0; first < total; first += st.chunk) {
        if (c->cancel.load()) {
            // with several ranks the peers are (or will be) waiting in the film reduce: take part in it before leaving, so that a
            // cancel on one rank cannot hang the others (pltb_cancel should still be called on every rank)
            if ((prm->flags & PLTB_RENDER_REDUCE) && c->world > 1)
                g_nccl.Reduce(c->accum, c->accum, n_acc, /* ncclFloat32 */ 7, /* ncclSum */ 0, 0, c->comm, s);
            cudaStreamSynchronize(s);
            return fail(PLTB_ERR_CANCELLED, "render cancelled");
        }
        NvtxRange pass_range("pltb:pass");
        pp.first = first;
        pp.count = (uint32_t) std::min<uint64_t>(st.chunk, total - first);
        CU(cudaMemsetAsync(st.counts, 0, kCountsTotal * sizeof(uint32_t), s));
        PROF();
        if (profile) KHook::mark(-1, c);
        // sample phase as dense per-bounce wavefronts: raygen + trace(0), then shade(i) / trace(i + 1)
        nvtxRangePushA("pltb:sample_phase");
        launch_raygen(dev_scene, st, pp, s);
        // one k_shade / k_nee_post launch per material class the scene contains (roughgrating first: the longest kernel)
        const uint32_t cm = scene->class_mask;
        const int n_cls = __builtin_popcount(cm), has_grating = (cm >> PLTB_BSDF_ROUGHGRATING) & 1;
        for (uint32_t b = 0; b < prm->max_depth; ++b) {
            for (int cls = kClasses - 1; cls >= 0; --cls)
                if (cm >> cls & 1u) launch_shade(dev_scene, st, b, cls, wf_grid, s);
            launches += n_cls;
            if (b + 1 < prm->max_depth) { launch_trace(dev_scene, st, b + 1, wf_grid, s); launches += 2; }
        }
        nvtxRangePop();
        PROF();
        nvtxRangePushA("pltb:solve_nee");
        const uint32_t nm = cm & kSmoothClassMask;   // the classes whose vertices can queue NEE items
        const int n_nee = __builtin_popcount(nm);
        launch_nee_light(dev_scene, st, nm, nee_grid, s);
        for (int cls = kClasses - 1; cls >= 0; --cls)
            if ((nm >> cls & 1u) && cls != PLTB_BSDF_ROUGHGRATING) launch_nee_post(dev_scene, st, cls, nee_grid, s);
        PROF();
        if (has_grating) launch_nee_post(dev_scene, st, PLTB_BSDF_ROUGHGRATING, nee_grid, s);
        nvtxRangePop();
        PROF();
        NvtxRange film_range("pltb:gather_splat");
        launch_splat(dev_scene, st, pp, c->accum, do_splat, d_wl, d_L, d_rays, d_stokes, stokes ? 1 : 0, dcap, s);
        PROF();
        // raygen, k_nee_light per class, shadow trace, k_nee_post per class (+ k_nee_eval for the grating), gather (+ splat)
        launches += 1 + n_nee + 1 + n_nee + (has_grating && !polarized && !rgb ? 2 : 0) + 1 + (do_splat ? (stokes && strip_splat ? 4 : 1) : 0);   // grating: + k_nee_scan, k_nee_eval; Stokes film: one strip launch per float4 group
    }
    CU(cudaGetLastError());
    PROF();   // tail event 0
    if ((prm->flags & PLTB_RENDER_REDUCE) && c->world > 1) {
        NvtxRange reduce_range("pltb:film_reduce");
        // one sum-reduce of the raw RGBW accumulator to rank 0 per render over NVLink
        int r = g_nccl.Reduce(c->accum, c->accum, n_acc, /* ncclFloat32 */ 7, /* ncclSum */ 0, 0, c->comm, s);
        if (r != 0) return fail(PLTB_ERR_CUDA, std::string("ncclReduce: ") + (g_nccl.GetErrorString ? g_nccl.GetErrorString(r) : "error"));
    }
    PROF();   // tail event 1: after the reduce
    const bool root = !(prm->flags & PLTB_RENDER_REDUCE) || c->rank == 0;
    if (film_out && root) {
        const float *src = c->accum; uint64_t n_out = n_acc;
        if (prm->flags & PLTB_RENDER_DEVELOP) {
            const uint64_t n_rgb = (uint64_t) W * H * (nch - 1);
            if (c->rgb_floats != n_rgb) {
                cudaFree(c->rgb); c->rgb = nullptr;
                CU(cudaMalloc(&c->rgb, n_rgb * sizeof(float)));
                c->rgb_floats = n_rgb;
            }
            launch_develop(c->accum, c->rgb, (uint64_t) W * H, nch, s);
            launches++;
            src = c->rgb; n_out = n_rgb;
        }
        PROF();   // tail event 2: after develop
        CU(cudaMemcpyAsync(film_out, src, n_out * sizeof(float),
                           (prm->flags & PLTB_RENDER_DEVICE_OUT) ? cudaMemcpyDeviceToDevice : cudaMemcpyDeviceToHost, s));
    }
    if (dcap) {
        if (dump->wavelengths) CU(cudaMemcpyAsync(dump->wavelengths, d_wl, dcap * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
        if (dump->radiance) CU(cudaMemcpyAsync(dump->radiance, d_L, dcap * 4 * sizeof(float), cudaMemcpyDeviceToHost, s));
        if (dump->n_rays) CU(cudaMemcpyAsync(dump->n_rays, d_rays, dcap * 2 * sizeof(uint32_t), cudaMemcpyDeviceToHost, s));
        if (d_stokes) CU(cudaMemcpyAsync(dump->stokes, d_stokes, dcap * 16 * sizeof(float), cudaMemcpyDeviceToHost, s));
    }
    if (!(film_out && root)) PROF();
    PROF();   // tail event 3: after the copies
    CU(cudaEventRecord(c->ev[1], s));
    if (stats) CU(cudaMemcpyAsync(c->h_counters, c->counters, 8 * sizeof(unsigned long long), cudaMemcpyDeviceToHost, s));
    CU(cudaStreamSynchronize(s));
    CU(cudaGetLastError());
    if (stats) {
        memset(stats, 0, sizeof(*stats));
        const unsigned long long *h = c->h_counters;
        stats->samples = total; stats->closest_rays = h[0]; stats->shadow_rays = h[1];
        stats->vertices = h[2]; stats->grating_vertices = h[3];
        stats->nee_other = h[4]; stats->nee_grating = h[5];
        stats->passes = n_pass;
        stats->nee_grating_eval = h[6];
        stats->kernel_launches = launches;
        CU(cudaEventElapsedTime(&stats->ms_total, c->ev[0], c->ev[1]));
        if (profile) {
            float ms;
            for (uint64_t p = 0; p < n_pass; ++p) {
                cudaEvent_t *e = &c->prof[5 * p];
                CU(cudaEventElapsedTime(&ms, e[0], e[1])); stats->ms_sample += ms;
                CU(cudaEventElapsedTime(&ms, e[1], e[2])); stats->ms_nee_other += ms;     // light + shadow rays + post<0>
                CU(cudaEventElapsedTime(&ms, e[2], e[3])); stats->ms_nee_grating += ms;   // post<1> + eval
                CU(cudaEventElapsedTime(&ms, e[3], e[4])); stats->ms_splat += ms;
            }
            for (size_t i = 1; i < c->kprof_n; ++i) {
                const int k = c->kprof_id[i];
                if (k < 0 || k >= PLTB_K_COUNT) continue;
                CU(cudaEventElapsedTime(&ms, c->kprof[i - 1], c->kprof[i]));
                stats->ms_kernel[k] += ms; stats->launches_kernel[k]++;
            }
            cudaEvent_t *t = &c->prof[5 * n_pass];
            CU(cudaEventElapsedTime(&stats->ms_reduce, t[0], t[1]));
            CU(cudaEventElapsedTime(&stats->ms_develop, t[1], t[2]));
            CU(cudaEventElapsedTime(&stats->ms_copy, t[2], t[3]));
        }
    }
    return 0;
}

}  // extern "C"
