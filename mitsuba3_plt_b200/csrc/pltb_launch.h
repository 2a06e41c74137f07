// pltb_launch.h -- launcher prototypes shared by pltb_kernels.cu and pltb_capi.cu
#pragma once
#include <cuda_runtime.h>

#include "pltb_integrator.cuh"

namespace pltb {
void launch_raygen(const DeviceScene &sc, const PassState &st, const PassParams &pp, cudaStream_t s);
void launch_trace(const DeviceScene &sc, const PassState &st, uint32_t bounce, int grid, cudaStream_t s);
void launch_shade(const DeviceScene &sc, const PassState &st, uint32_t bounce, int cls, int grid, cudaStream_t s);
void launch_nee_light(const DeviceScene &sc, const PassState &st, int grid, cudaStream_t s);
void launch_nee_post(const DeviceScene &sc, const PassState &st, int cls, int grid, cudaStream_t s);
void launch_splat(const DeviceScene &sc, const PassState &st, const PassParams &pp, float *accum, int do_splat, float *dump_wl,
                  float *dump_L, uint64_t dump_cap, cudaStream_t s);
void launch_develop(const float *accum, float *rgb, uint64_t n_pixels, cudaStream_t s);
}  // namespace pltb
