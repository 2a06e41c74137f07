// pltb_kernels.cu -- sm_100a kernels of the plt render path and their launchers.
//   k_raygen : ray generation + bounce-0 closest hit, one thread per Monte-Carlo sample
//   k_shade<c>(i) / k_trace(i) : the sample-phase loop unrolled over bounces as dense wavefronts (grid-stride over device-side
//              work lists; one material class per k_shade launch)
//   k_nee<c> : next-event estimation, grid-stride over the NEE work queues filled by k_shade
//   k_splat  : stage 3 (ordered accumulation, spectrum -> sRGB, film splat with vector atomics)
//   k_develop: HDRFilm::develop weight division
// Path state lives in SoA arrays indexed depth-major (pltb_integrator.cuh) so warp accesses are coalesced float4s; queue
// appends use warp-aggregated atomics (match_any + popc + one atomicAdd per group).
#include <cuda_runtime.h>

#include "pltb_stages.cuh"
#include "pltb_launch.h"

namespace pltb {

// Warp-aggregated list append: lanes appending to the same list elect a leader that reserves the slots with one atomicAdd.
struct DevPush {
    __device__ __forceinline__ void operator()(uint32_t *list, uint32_t *count, uint32_t entry) const {
        const unsigned active = __activemask();
        const unsigned peers = __match_any_sync(active, (unsigned long long) count);
        const int lane = threadIdx.x & 31, leader = __ffs(peers) - 1;
        uint32_t base = 0;
        if (lane == leader) base = atomicAdd(count, (uint32_t) __popc(peers));
        base = __shfl_sync(peers, base, leader);
        list[base + __popc(peers & ((1u << lane) - 1u))] = entry;
    }
};

struct DevFilm {
    float *accum;   // RGBW, 4 floats per pixel
    __device__ __forceinline__ void add(size_t pixel, const float values[4], float w) {
        float4 v = make_float4(values[0] * w, values[1] * w, values[2] * w, values[3] * w);
        atomicAdd(reinterpret_cast<float4 *>(accum) + pixel, v);   // red.global.add.v4.f32 (sm_90+)
    }
};

__device__ __forceinline__ void flush_counters(const uint32_t cnt[4], unsigned long long *counters) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t v = cnt[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(counters + k, (unsigned long long) v);
    }
}

__global__ void __launch_bounds__(128) k_raygen(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                const __grid_constant__ PassParams pp) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    if (local < pp.count) {
        DevPush push;
        stage_raygen(sc, st, pp, local, push, cnt);
    }
    flush_counters(cnt, st.counters);
}

// grid-stride over a device-side work list whose length is only known on the device
__global__ void __launch_bounds__(128) k_trace(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               uint32_t bounce) {
    const uint32_t n = st.counts[4 * bounce];
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    DevPush push;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_trace(sc, st, st.live[e], bounce, push, cnt);
    flush_counters(cnt, st.counters);
}

template <int Class>
__global__ void __launch_bounds__(128) k_shade(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               uint32_t bounce) {
    const uint32_t n = st.counts[4 * bounce + 1 + Class];
    const uint32_t *list = st.shade_q[Class];
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    DevPush push;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_shade<Class>(sc, st, list[e], bounce, push, cnt);
    flush_counters(cnt, st.counters);
}

template <int Class>
__global__ void __launch_bounds__(128) k_nee(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st) {
    const uint32_t n = st.counts[kCountsNee + Class];
    const uint32_t *queue = st.queue[Class];
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(st.counters + 4 + Class, (unsigned long long) n);   // queue sizes
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_nee<Class>(sc, st, queue[e], cnt);
    flush_counters(cnt, st.counters);
}

__global__ void __launch_bounds__(256) k_splat(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               const __grid_constant__ PassParams pp, float *accum, int do_splat,
                                               float *dump_wl, float *dump_L, uint64_t dump_cap) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= pp.count) return;
    DevFilm film{ accum };
    S4 L = stage_splat(sc, st, pp, local, film, do_splat != 0);
    uint32_t pix, s_rel;
    split_work(pp, pp.first + local, pix, s_rel);
    const uint64_t di = (uint64_t) pix * pp.spp_count + s_rel;   // dump order: pixel-major, then sample (include/pltb.h)
    if (di < dump_cap) {
        if (dump_L) reinterpret_cast<float4 *>(dump_L)[di] = s4_to_f4(L);
        if (dump_wl) reinterpret_cast<float4 *>(dump_wl)[di] = st.wl[local];
    }
}

// HDRFilm::develop (src/films/hdrfilm.cpp:304-400): rgb / (w == 0 ? 1 : w)
__global__ void k_develop(const float4 *accum, float *rgb, uint64_t n_pixels) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    float4 a = accum[i];
    float d = a.w == 0.f ? 1.f : a.w;
    rgb[3 * i] = a.x / d; rgb[3 * i + 1] = a.y / d; rgb[3 * i + 2] = a.z / d;
}

void launch_raygen(const DeviceScene &sc, const PassState &st, const PassParams &pp, cudaStream_t s) {
    k_raygen<<<(pp.count + 127) / 128, 128, 0, s>>>(sc, st, pp);
}
void launch_trace(const DeviceScene &sc, const PassState &st, uint32_t bounce, int grid, cudaStream_t s) {
    k_trace<<<grid, 128, 0, s>>>(sc, st, bounce);
}
void launch_shade(const DeviceScene &sc, const PassState &st, uint32_t bounce, int cls, int grid, cudaStream_t s) {
    if (cls) k_shade<1><<<grid, 128, 0, s>>>(sc, st, bounce);
    else k_shade<0><<<grid, 128, 0, s>>>(sc, st, bounce);
}
void launch_nee(const DeviceScene &sc, const PassState &st, int cls, int grid, cudaStream_t s) {
    if (cls) k_nee<1><<<grid, 128, 0, s>>>(sc, st);
    else k_nee<0><<<grid, 128, 0, s>>>(sc, st);
}
void launch_splat(const DeviceScene &sc, const PassState &st, const PassParams &pp, float *accum, int do_splat, float *dump_wl,
                  float *dump_L, uint64_t dump_cap, cudaStream_t s) {
    k_splat<<<(pp.count + 255) / 256, 256, 0, s>>>(sc, st, pp, accum, do_splat, dump_wl, dump_L, dump_cap);
}
void launch_develop(const float *accum, float *rgb, uint64_t n_pixels, cudaStream_t s) {
    k_develop<<<(unsigned) ((n_pixels + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4 *>(accum), rgb, n_pixels);
}

}  // namespace pltb
