// pltb_launch.h -- launcher prototypes shared by pltb_kernels.cu and pltb_capi.cu
#pragma once
#include <cuda_runtime.h>

#include "pltb_integrator.cuh"

namespace pltb {
// Optional per-kernel hook (PLTB_RENDER_PROFILE): called by the launchers after every kernel launch with its PLTB_K_* id.
struct LaunchHook { void (*fn)(int kernel, void *user) = nullptr; void *user = nullptr; };
extern thread_local LaunchHook g_launch_hook;
void launch_raygen(const DeviceScene &sc, const PassState &st, const PassParams &pp, cudaStream_t s);
void launch_trace(const DeviceScene &sc, const PassState &st, uint32_t bounce, int grid, cudaStream_t s);
void launch_shade(const DeviceScene &sc, const PassState &st, uint32_t bounce, int cls, int grid, cudaStream_t s);
void launch_nee_light(const DeviceScene &sc, const PassState &st, uint32_t class_mask, int grid, cudaStream_t s);
void launch_nee_post(const DeviceScene &sc, const PassState &st, int cls, int grid, cudaStream_t s);
void launch_splat(const DeviceScene &sc, const PassState &st, const PassParams &pp, float *accum, int do_splat, float *dump_wl,
                  float *dump_L, uint32_t *dump_rays, float *dump_stokes, int stokes, uint64_t dump_cap, cudaStream_t s);
void launch_develop(const float *accum, float *rgb, uint64_t n_pixels, int nch, cudaStream_t s);
}  // namespace pltb
