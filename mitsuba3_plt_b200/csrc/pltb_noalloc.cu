// pltb_noalloc.cu -- pltb_film_develop_noalloc (include/pltb.h): HDRFilm::develop of a caller-supplied raw film accumulator with
// NO device allocation.  No counterpart in the reference; added for callers that cannot
// call cudaMalloc (CUDA graph capture, tight serving loops, harnesses that forbid it).
//
// pltb_render's develop step keeps the developed image in a cudaMalloc'ed buffer (pltb_context::rgb) until it is copied out.
// Here both the staged accumulator and the developed image live in ONE file-scope __device__ array of fixed size; every GPU
// that loads the library holds it (2 * PLTB_NOALLOC_MAX_FLOATS floats = 32 MiB) until the process exits.
//
// The array is shared by every call on a device and keeps the previous call's contents.  Nothing has to start at zero: a call
// first overwrites the n floats of the input region it will read, and the kernel writes every float of the output region that is
// copied out.  Calls that overlap on the GPU (different streams, or PLTB_DEVELOP_ASYNC calls not yet complete) RACE on the array.
#include <cuda_runtime.h>

#include <mutex>
#include <string>

#include "../../include/pltb.h"

namespace pltb {
int set_last_error(int code, const char *msg);   // pltb_capi.cu: the thread's pltb_last_error() text
}

// First half: the staged accumulator; second half: the developed image.  In HOST code this name is only a stand-in for the
// symbol, not a device address -- host code goes through cudaGetSymbolAddress (scratch_for_current_device below).
static __device__ __align__(256) float g_noalloc_scratch[2 * (size_t) PLTB_NOALLOC_MAX_FLOATS];

namespace {

constexpr size_t kMaxFloats = PLTB_NOALLOC_MAX_FLOATS;
constexpr int kMaxDevices = 64;

std::mutex g_lock;
float *g_scratch_ptr[kMaxDevices];   // device address of g_noalloc_scratch, per device ordinal

// HDRFilm::develop (src/films/hdrfilm.cpp:304-400) for nch accumulator channels [R G B W aov...]: every channel but W divided by
// (w == 0 ? 1 : w), same arithmetic as k_develop / k_develop_n (pltb_kernels.cu).
__global__ void __launch_bounds__(256) k_develop_static(const float *__restrict__ accum, float *__restrict__ out, uint64_t n_pixels, int nch) {
    for (uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x; i < n_pixels; i += (uint64_t) gridDim.x * blockDim.x) {
        const float w = accum[nch * i + 3], d = w == 0.f ? 1.f : w;
        int o = 0;
        for (int c = 0; c < nch; ++c)
            if (c != 3) out[(nch - 1) * i + o++] = accum[nch * i + c] / d;
    }
}

int cuda_fail(const char *what, cudaError_t e) {
    return pltb::set_last_error(PLTB_ERR_CUDA, (std::string("pltb_film_develop_noalloc: ") + what + ": " + cudaGetErrorString(e)).c_str());
}

// once per device, with that device current
int scratch_for_current_device(float **out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return cuda_fail("cudaGetDevice (the plt path has no CPU fallback)", e);
    if (dev < 0 || dev >= kMaxDevices) return pltb::set_last_error(PLTB_ERR_UNSUPPORTED, "pltb_film_develop_noalloc: device ordinal >= 64");
    std::lock_guard<std::mutex> guard(g_lock);
    if (!g_scratch_ptr[dev]) {
        float *p = nullptr;
        e = cudaGetSymbolAddress((void **) &p, g_noalloc_scratch);
        if (e != cudaSuccess || !p) return cuda_fail("cudaGetSymbolAddress(g_noalloc_scratch)", e);
        g_scratch_ptr[dev] = p;
    }
    *out = g_scratch_ptr[dev];
    return 0;
}

}  // namespace

extern "C" int pltb_film_develop_noalloc(const float *accum, uint32_t width, uint32_t height, uint32_t channels, uint32_t flags,
                                         float *out, void *stream) {
    if (!accum || !out) return pltb::set_last_error(PLTB_ERR_INVALID, "pltb_film_develop_noalloc: null accum / out");
    if (channels != 4 && channels != 16)
        return pltb::set_last_error(PLTB_ERR_INVALID, "pltb_film_develop_noalloc: channels must be 4 (R G B W) or 16 (R G B W + 12 Stokes AOVs)");
    if (flags & ~(uint32_t) PLTB_DEVELOP_ASYNC) return pltb::set_last_error(PLTB_ERR_INVALID, "pltb_film_develop_noalloc: unknown flag");
    const uint64_t n_pixels = (uint64_t) width * height, n_in = n_pixels * channels, n_out = n_pixels * (channels - 1);
    if (n_pixels == 0) return 0;
    if (n_in > kMaxFloats)
        return pltb::set_last_error(PLTB_ERR_NOMEM, ("pltb_film_develop_noalloc: width * height * channels = " + std::to_string(n_in) +
                                                     " floats exceeds PLTB_NOALLOC_MAX_FLOATS = " + std::to_string(kMaxFloats) +
                                                     " (the static scratch array; use pltb_render's develop for larger films)").c_str());
    float *scratch = nullptr;
    if (int rc = scratch_for_current_device(&scratch)) return rc;
    cudaStream_t s = (cudaStream_t) stream;
    float *d_in = scratch, *d_out = scratch + kMaxFloats;
    cudaError_t e = cudaMemcpyAsync(d_in, accum, n_in * sizeof(float), cudaMemcpyDefault, s);   // host or device source (UVA)
    if (e != cudaSuccess) return cuda_fail("copy in", e);
    const uint64_t blocks = (n_pixels + 255) / 256;
    const unsigned grid = (unsigned) (blocks < 148u * 8u ? blocks : 148u * 8u);   // grid-stride beyond 8 blocks per SM
    k_develop_static<<<grid, 256, 0, s>>>(d_in, d_out, n_pixels, (int) channels);
    if ((e = cudaGetLastError()) != cudaSuccess) return cuda_fail("k_develop_static launch", e);
    if ((e = cudaMemcpyAsync(out, d_out, n_out * sizeof(float), cudaMemcpyDefault, s)) != cudaSuccess) return cuda_fail("copy out", e);
    if (!(flags & PLTB_DEVELOP_ASYNC) && (e = cudaStreamSynchronize(s)) != cudaSuccess) return cuda_fail("synchronize", e);
    return 0;
}
