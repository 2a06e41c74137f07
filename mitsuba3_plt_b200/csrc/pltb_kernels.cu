// pltb_kernels.cu -- sm_100a kernels of the plt render path and their launchers.
//   k_raygen : ray generation + bounce-0 closest hit, one thread per Monte-Carlo sample
//   k_shade<c>(i) / k_trace(i) : the sample-phase loop unrolled over bounces as dense wavefronts (grid-stride over device-side
//              work lists; one material class per k_shade launch)
//   k_nee<c> : next-event estimation, grid-stride over the record-slot runs of a material class
//   k_splat  : stage 3 (ordered accumulation, spectrum -> sRGB, film splat with vector atomics)
//   k_develop: HDRFilm::develop weight division
// Path state lives in SoA arrays indexed depth-major (pltb_integrator.cuh) so warp accesses are coalesced float4s.  k_shade
// appends nothing with atomics: record slots and next-bounce list positions follow from the position in the shade list
// (warp_slot_prefix); the routing and NEE lists use warp-aggregated atomics (match_any + popc + one atomicAdd per group).
#include <cuda_runtime.h>
#include <cstdlib>
#include <type_traits>

#include "pltb_stages_pol.cuh"
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
// one float4 group G of a 16-channel Stokes film (R G B W | S0.RGB S1.R | S1.GB S2.RG | S2.B S3.RGB)
template <int G> struct DevFilmGroup {
    float *accum;
    __device__ __forceinline__ void add(size_t pixel, const float values[4], float w) {
        float4 v = make_float4(values[0] * w, values[1] * w, values[2] * w, values[3] * w);
        atomicAdd(reinterpret_cast<float4 *>(accum) + 4 * pixel + G, v);
    }
};

__device__ __forceinline__ void prefetch_l2(const void *p) { asm volatile("prefetch.global.L2 [%0];" ::"l"(p)); }

__device__ __forceinline__ void flush_counters(const uint32_t cnt[4], unsigned long long *counters) {
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        uint32_t v = cnt[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
        if ((threadIdx.x & 31) == 0 && v) atomicAdd(counters + k, (unsigned long long) v);
    }
}

// Record slots (PassState): the vertices shaded at bounce b by class c own the run [slot_base(b, c), + counts[cnt_shade(b, c)]),
// the runs laid out in launch order (bounce-major, then class).  One warp computes the exclusive prefix over the (b, c) pairs
// of bounces [0, nb) -- kClasses * nb <= 112 shade-queue counts, every lane a few loads and a shuffle scan -- and hands each pair
// to `f(b, c, base, count)` (called by the lane that holds the pair).
template <class F> __device__ __forceinline__ void warp_slot_prefix(const PassState &st, uint32_t nb, F &&f) {
    const unsigned lane = threadIdx.x & 31u;
    uint32_t carry = 0;
    for (uint32_t k0 = 0; k0 < nb * kClasses; k0 += 32u) {
        const uint32_t k = k0 + lane, b = k / kClasses, c = k % kClasses;
        const uint32_t v = k < nb * kClasses ? st.counts[cnt_shade(b, c)] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((int) lane >= o) incl += t;
        }
        if (k < nb * kClasses) f(b, c, carry + incl - v, v);
        carry += __shfl_sync(0xffffffffu, incl, 31);
    }
}
// slot_base(bounce, Class) for a k_shade launch and the first position of the class in the next bounce's live list, block-wide
// (contains a barrier)
__device__ __forceinline__ void block_slot_base(const PassState &st, uint32_t bounce, uint32_t cls, uint32_t &slot_base, uint32_t &live_base) {
    __shared__ uint32_t s_base[2];
    if (threadIdx.x < 32)
        warp_slot_prefix(st, bounce + 1, [&](uint32_t b, uint32_t c, uint32_t base, uint32_t) {
            if (b == bounce && c == cls) s_base[0] = base;
            if (b == bounce && c == 0) s_base[1] = base;
        });
    __syncthreads();
    slot_base = s_base[0]; live_base = s_base[0] - s_base[1];
}
// number of positions of the live list of bounce >= 1: the vertices shaded at the bounce before (PassState::live)
__device__ __forceinline__ uint32_t live_count(const PassState &st, uint32_t bounce) {
    uint32_t n = 0;
#pragma unroll
    for (uint32_t c = 0; c < kClasses; ++c) n += st.counts[cnt_shade(bounce - 1, c)];
    return n;
}
// The NEE work list of one class: its slot runs, bounce by bounce.  start[b] = number of list positions before bounce b.
struct ClassRuns {
    uint32_t start[PLTB_MAX_DEPTH + 1], base[PLTB_MAX_DEPTH];
    __device__ __forceinline__ uint32_t slot(uint32_t x) const {
        uint32_t b = 0;
        while (x >= start[b + 1]) ++b;
        return base[b] + (x - start[b]);
    }
};
__device__ __forceinline__ uint32_t block_class_runs(const PassState &st, uint32_t cls, ClassRuns &runs) {   // contains a barrier; returns the list size
    if (threadIdx.x < 32) {
        const unsigned lane = threadIdx.x;
        warp_slot_prefix(st, st.max_depth, [&](uint32_t b, uint32_t c, uint32_t base, uint32_t count) {
            if (c == cls) { runs.base[b] = base; runs.start[b + 1] = count; }
        });
        __syncwarp();
        // start[] = prefix of the per-bounce counts: a warp scan (lane b holds bounce b; PLTB_MAX_DEPTH <= 32)
        const uint32_t v = lane < st.max_depth ? runs.start[lane + 1] : 0u;
        uint32_t incl = v;
#pragma unroll
        for (int o = 1; o < 32; o <<= 1) {
            const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
            if ((int) lane >= o) incl += t;
        }
        if (lane < st.max_depth) runs.start[lane + 1] = incl;
        if (lane == 0) runs.start[0] = 0u;
    }
    __syncthreads();
    return runs.start[st.max_depth];
}

// (register budgets measured on cbox-grating, ms per 2^30 samples: k_raygen 72 regs 75.4, 64 regs 73.4, 48 regs 73.3; k_classify 80 regs 28.5,
// 64 regs 29.4, 40 regs 30.2; k_gather 60 regs 18.0, 40 regs 20.2, 32 regs 32.0; k_nee_light 64 regs 27-28, 48 regs 30.8, 40 regs 33.6)
#ifndef PLTB_RAYGEN_THREADS
#define PLTB_RAYGEN_THREADS 128
#endif
// a list append that does nothing: k_raygen appends the hits of a whole block at once
struct NoPush { __device__ __forceinline__ void operator()(uint32_t *, uint32_t *, uint32_t) const {} };
// The block's hits go to the shade queues as ONE run per class (one atomicAdd per class and block): a k_shade warp then reads the
// path state of consecutive samples -- appended warp by warp, the 16 or so entries a warp contributes to a class sat between
// those of unrelated warps.  No warp waits for another: each leaves the classes of its lanes in shared memory and the LAST one to
// arrive appends for the whole block (with barriers, 17 % of the kernel's stall samples sat at the first one: a block waited for
// its slowest traversal).
__global__ void __launch_bounds__(PLTB_RAYGEN_THREADS, 1024 / PLTB_RAYGEN_THREADS) k_raygen(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                const __grid_constant__ PassParams pp) {
    constexpr uint32_t kWarps = PLTB_RAYGEN_THREADS / 32;
    __shared__ uint32_t s_cnt[kWarps][kClasses + 1];
    __shared__ uint32_t s_base[kClasses];
    __shared__ uint32_t s_arrived;
    __shared__ uint8_t s_cls[PLTB_RAYGEN_THREADS];
    const uint32_t block_first = blockIdx.x * blockDim.x, local = block_first + threadIdx.x, lane = threadIdx.x & 31u;
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    if (threadIdx.x == 0) s_arrived = 0u;
    __syncthreads();   // (all warps are still together here)
    int cls = -1;
    if (local < pp.count) {
        NoPush push;
        cls = stage_raygen(sc, st, pp, local, push, cnt);
    }
    flush_counters(cnt, st.counters);
    s_cls[threadIdx.x] = cls >= 0 ? (uint8_t) cls : (uint8_t) 0xFF;
    __threadfence_block();
    __syncwarp();
    uint32_t order = 0;
    if (lane == 0) order = atomicAdd(&s_arrived, 1u);
    order = __shfl_sync(0xffffffffu, order, 0);
    if (order != kWarps - 1) return;
    __threadfence_block();
    // the last warp of the block: group g = the lanes of warp g
    for (uint32_t k = lane; k < kWarps * (kClasses + 1); k += 32u) (&s_cnt[0][0])[k] = 0u;
    __syncwarp();
    uint32_t rank[kWarps], c[kWarps];
#pragma unroll
    for (uint32_t g = 0; g < kWarps; ++g) {
        c[g] = s_cls[32u * g + lane];
        const unsigned peers = __match_any_sync(0xffffffffu, c[g] != 0xFFu ? c[g] : 0x100u + lane);
        rank[g] = __popc(peers & ((1u << lane) - 1u));
        if (c[g] != 0xFFu && rank[g] == 0) s_cnt[g][c[g]] = (uint32_t) __popc(peers);
    }
    __syncwarp();
    if (lane < (uint32_t) kClasses) {
        uint32_t off = 0;
#pragma unroll
        for (uint32_t g = 0; g < kWarps; ++g) { const uint32_t v = s_cnt[g][lane]; s_cnt[g][lane] = off; off += v; }
        if (off) s_base[lane] = atomicAdd(st.counts + cnt_shade(0, lane), off);
    }
    __syncwarp();
#pragma unroll
    for (uint32_t g = 0; g < kWarps; ++g)
        if (c[g] != 0xFFu) st.shade_q[c[g]][s_base[c[g]] + s_cnt[g][c[g]] + rank[g]] = block_first + 32u * g + lane;
}

// Closest hit for the live paths of bounce >= 1.  Secondary rays are incoherent and their traversal lengths differ by orders
// of magnitude (most leave the open box after a few nodes, some walk a 960-triangle sphere), so a ray-per-thread loop leaves
// two thirds of the lanes idle.  Persistent warps instead: every lane owns a resumable Traversal; whenever lanes finish, the
// warp reserves new rays for them with one atomicAdd on the bounce's cursor, then all lanes advance by a bounded number of
// inner-node steps followed by one leaf step (while-while order).
// (Measured and rejected: feeding the lanes through a three-stage prefetch pipeline -- warp-private pools of list positions,
// entry and ray loads issued one round ahead -- made this kernel 1.5x slower (0.89 -> 1.37 ms per pass) at equal results: the
// DRAM round trip of a refill is already covered by the other 20+ resident warps, while the extra per-round bookkeeping and
// the loss of globally consecutive ray batches are not.)
// Shadow = true: the same loop over the NEE shadow-ray list (any-hit; an occluded vertex gets ds.pdf = 0 in its record).
#ifndef PLTB_TRACE_INNER
#define PLTB_TRACE_INNER 6   // inner-node steps per round before the leaf step
#endif
#ifndef PLTB_TRACE_MINB
#define PLTB_TRACE_MINB 8
#endif
#ifndef PLTB_TRACE_MINB_SHADOW
#define PLTB_TRACE_MINB_SHADOW 8
#endif
// register budgets: 64 for both forms (8 blocks of 128 threads per SM).  With the traversal state in registers (the stack array
// outside the Traversal struct) k_trace 95.7 -> 81.3, k_trace<shadow> 72.3 -> 61.2 ms per 2^30 samples on cbox-grating; closest hit at
// 80 / 48 registers 89.7 / 88.2, any hit at 48 / 40 registers 63.7 / 75.9 (profiles/r02_ab_session17.txt).
// Measured and rejected on top of that: stack entries 0..7 / 0..11 in shared memory (82.6, any hit 67.0); inner steps while at
// least as many lanes want one as wait at a leaf (83.7 / 66.8); 4 or 8 inner steps per round (86.6 / 84.1); reserving 64 or 128
// list positions per atomicAdd and handing them out by shuffle (82.5 / 87.8); requesting a leaf's next primitive before the
// current one is tested (85.8).
template <bool Shadow>
__global__ void __launch_bounds__(128, Shadow ? PLTB_TRACE_MINB_SHADOW : PLTB_TRACE_MINB) k_trace(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               uint32_t bounce) {
    const uint32_t n = Shadow ? st.counts[kCntShadow] : live_count(st, bounce);
    uint32_t *cursor = Shadow ? st.counts + kCntShadowCursor : st.counts + cnt_cursor(bounce);
    const unsigned lane = threadIdx.x & 31u, lt_mask = (1u << lane) - 1u;
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    unsigned long long stack_mem[kBvhStack];
    Traversal tr;
    tr.done = true;
    bool active = false, exhausted = false;
    size_t local = 0;   // path id (closest hit) or vertex record index (shadow)
    uint32_t epos = 0;  // position of the lane's ray in the live list
    for (;;) {
        __syncwarp();
        const unsigned need = __ballot_sync(0xffffffffu, !active);
        if (need && !exhausted) {
            uint32_t base = 0;
            const int leader = __ffs(need) - 1;
            if ((int) lane == leader) base = atomicAdd(cursor, (uint32_t) __popc(need));
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!active) {
                const uint32_t e = base + __popc(need & lt_mask);
                if (e < n) {
                    Ray ray;
                    bool have = true;
                    if (Shadow) {
                        local = st.shadow_q[e];   // record slot
                        ray = nee_shadow_ray(st, local);
                    } else {
                        local = st.live[e]; epos = e;
                        have = local != kNoPath;   // else the path ended at the previous bounce: the lane asks again next round
                        if (have) {
                            float4 o = st.p_ray_o[local], d = st.p_ray_d[local];
                            ray.o = mk3(o.x, o.y, o.z); ray.d = mk3(d.x, d.y, d.z); ray.maxt = o.w;
                            cnt[0]++;
                        }
                    }
                    if (have) {
                        tr.begin(sc, ray, stack_mem);
                        active = true;
                    }
                }
            }
            exhausted = base + __popc(need) >= n;   // warp-uniform
        }
        if (__ballot_sync(0xffffffffu, active) == 0) {   // warp-uniform
            if (exhausted) break;
            continue;   // a refill that met ended paths only
        }
#pragma unroll 1
        for (int k = 0; k < PLTB_TRACE_INNER; ++k)
            if (active && !tr.done && tr.node >= 0) tr.inner_step(sc);
        if (active && !tr.done && tr.node < 0) tr.template leaf_step<Shadow>(sc);
        if (active && tr.done) {
            if (Shadow) {
                if (tr.best.shape >= 0) st.v_w[local].w = 0.f;   // occluded: ds.pdf = 0 (scene.cpp:330-340)
                active = false;
                continue;
            }
            // results are written in place; k_classify walks the live list in order afterwards so that the shade / NEE work
            // lists stay (nearly) sorted by sample id and the per-sample state accesses of later kernels stay coalesced
            if (tr.best.shape >= 0) {
                st.p_hit[local] = mkf4(tr.best.t, tr.best.u, tr.best.v, u2f(tr.best.prim | (tr.best.cls << kHitClsShift)));
                st.hit_cls[epos] = (uint8_t) tr.best.cls;
            } else {
                st.hit_cls[epos] = 0xFF;
            }
            active = false;
        }
    }
    flush_counters(cnt, st.counters);
}

// Routes the results of k_trace: hits to the shade queue of their material class, misses finalised in place.  k_trace leaves the
// class (or 0xFF for a miss) in hit_cls[position in the live list], so this pass reads two dense arrays (4 + 1 bytes per ray) and
// only the misses touch per-path state.  Appending per warp (one atomicAdd per 32 entries and class) made this pass 15 % faster
// than the block-wide form below, but left the queues as interleaved 10-30 entry pieces of unrelated warps: with one run per
// block and class k_shade reads the state of consecutive paths (k_shade<other> 60.9 -> 54.2 ms per 2^30 samples on cbox, 3102 ->
// 2944 ms per render on veach-mis; profiles/r02_ab_session17.txt).
__global__ void __launch_bounds__(256) k_classify(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                  uint32_t bounce) {
    // A block takes 1024 consecutive entries at a time (warp w the entries 128 w .. 128 w + 127, in four rounds of 32) and reserves
    // ONE run per class in the shade queues for all of them: the 32 (warp, round) groups count their hits per class, warp 0 scans
    // the groups, lanes 0..6 issue the (at most seven, concurrent) atomicAdds.  List order is kept within the block's entries.
    const uint32_t n = live_count(st, bounce);
    const uint32_t lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    constexpr uint32_t kPerBlock = 8u * 128u;
    __shared__ uint32_t s_cnt[32][kClasses + 1];   // [group = 4 * warp + round][class]: hits, then their offset in the block's run
    __shared__ uint32_t s_base[kClasses];
    DevPush push;
    for (uint32_t base = blockIdx.x * kPerBlock; base < n; base += gridDim.x * kPerBlock) {
        for (uint32_t k = threadIdx.x; k < 32u * (kClasses + 1); k += blockDim.x) (&s_cnt[0][0])[k] = 0u;
        __syncthreads();
        uint32_t local[4], cls[4], rank[4];
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const uint32_t e = base + 128u * warp + 32u * j + lane;
            local[j] = e < n ? st.live[e] : kNoPath; cls[j] = e < n ? (uint32_t) st.hit_cls[e] : 0xFFu;
        }
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            const bool hit = local[j] != kNoPath && cls[j] != 0xFFu;
            const unsigned peers = __match_any_sync(0xffffffffu, hit ? cls[j] : 0x100u + lane);
            rank[j] = __popc(peers & ((1u << lane) - 1u));
            if (hit && rank[j] == 0) s_cnt[4 * warp + j][cls[j] & 7u] = (uint32_t) __popc(peers);
        }
        __syncthreads();
        if (warp == 0) {
            uint32_t mine = 0;   // lane c: hits of class c in the block's entries
#pragma unroll
            for (uint32_t c = 0; c < (uint32_t) kClasses; ++c) {
                const uint32_t v = s_cnt[lane][c];
                uint32_t incl = v;
#pragma unroll
                for (int o = 1; o < 32; o <<= 1) {
                    const uint32_t t = __shfl_up_sync(0xffffffffu, incl, o);
                    if ((int) lane >= o) incl += t;
                }
                s_cnt[lane][c] = incl - v;
                const uint32_t total = __shfl_sync(0xffffffffu, incl, 31);
                if (lane == c) mine = total;
            }
            if (lane < (uint32_t) kClasses && mine) s_base[lane] = atomicAdd(st.counts + cnt_shade(bounce, lane), mine);
        }
        // the per-path record of the misses (rng state, rng increment, flags: one sector), requested for all four rounds at once
        // and before the barrier that waits for the block's atomics
        uint64_t m_state[4], m_inc[4], m_flags[4];
#pragma unroll
        for (int j = 0; j < 4; ++j)
            if (local[j] != kNoPath && cls[j] == 0xFFu) {
                m_state[j] = st.rng_state[local[j]]; m_inc[j] = st.rng_inc[local[j]]; m_flags[j] = st.p_flags[local[j]];
            }
        __syncthreads();
#pragma unroll
        for (int j = 0; j < 4; ++j) {
            if (local[j] == kNoPath) continue;
            if (cls[j] != 0xFFu) {
                const uint32_t c = cls[j] & 7u;
                st.shade_q[c][s_base[c] + s_cnt[4 * warp + j][c] + rank[j]] = local[j];
            } else if (sc.env < 0) {   // miss without environment: route_hit's miss branch on the values already loaded
                uint64_t state = m_state[j];
                for (int k = 0; k < 6; ++k) state = state * 0x5851f42d4c957f2dULL + m_inc[j];
                st.rng_state[local[j]] = state;
                st.v_mask[local[j]] = pack_mask(m_flags[j]);
                st.n_vertices[local[j]] = bounce + 1;
            } else {
                Hit h; h.t = 0.f; h.u = h.v = 0.f; h.prim = 0; h.cls = 0; h.shape = -1;
                PCG32 rng; rng.state = m_state[j]; rng.inc = m_inc[j];
                route_hit(sc, st, local[j], bounce, h, rng, false, push);   // miss: finalised in place
            }
        }
        __syncthreads();
    }
}

// (register budget: ptxas settles on 128 registers = 4 blocks/SM; forcing 3, 5, 6 or 8 blocks/SM measured 495 / 477 / 490 / 466 ms
// against 462 ms for the whole sample phase, profiles/r01_ab_shade_occupancy.txt)
#ifndef PLTB_SHADE_MINBLOCKS
#define PLTB_SHADE_MINBLOCKS 5   // blocks of 128 threads per SM the register allocation must allow: 5 = 96 registers (4: 1789, 5: 1804, 6: 1792 Msamples/s on C2)
#endif
template <int Class>
__global__ void __launch_bounds__(128, PLTB_SHADE_MINBLOCKS) k_shade(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               uint32_t bounce) {
    const uint32_t n = st.counts[cnt_shade(bounce, Class)];
    const uint32_t *list = st.shade_q[Class];
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    DevPush push;
    uint32_t slot_base, live_base;   // record slot / next live-list position of list position e: base + e
    block_slot_base(st, bounce, Class, slot_base, live_base);
    // (measured and rejected: pulling the next item's path state towards L2 while the current one is shaded)
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_shade<Class>(sc, st, list[e], bounce, slot_base + e, live_base + e, push, cnt);
    flush_counters(cnt, st.counters);
}

// NEE step 1 for the queue of one material class (the stage itself is class-independent)
__global__ void __launch_bounds__(256) k_nee_light(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                   int cls) {
    __shared__ ClassRuns runs;
    const uint32_t n = block_class_runs(st, (uint32_t) cls, runs);
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    uint32_t items = 0;   // queue sizes: roughgrating / everything else
    DevPush push;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t slot = runs.slot(e);
        if (st.rec_entry[slot] == kNoRecord) continue;
        items++;
        stage_nee_light(sc, st, slot, push, cnt);
    }
    __syncwarp();
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) items += __shfl_xor_sync(0xffffffffu, items, o);
    if ((threadIdx.x & 31) == 0 && items) atomicAdd(st.counters + (cls == PLTB_BSDF_ROUGHGRATING ? 5 : 4), (unsigned long long) items);
    flush_counters(cnt, st.counters);
}

// solve_replay_NEE's BSDF half per material class.  roughgrating (not Rgb): this kernel finishes the unlit / malformed items and
// lists the lit ones in the scan list (stage_nee_post<kNeeFilter>); k_nee_scan walks that list densely, k_nee_eval the items inside a window.
// (Measured before: scan inside this kernel, one item per thread 53.6 ms per 2^30 samples on cbox-grating; block-level hand-over
// of the lit items to dense threads 47.2 ms with 29 % of the warp samples at its barrier; this form: see profiles/r02_ab_session16.txt)
template <int Class, bool Rgb>
__global__ void __launch_bounds__(128) k_nee_post(const __grid_constant__ DeviceScene sc,
                                                                      const __grid_constant__ PassState st) {
    __shared__ ClassRuns runs;
    const uint32_t n = block_class_runs(st, (uint32_t) Class, runs);
    DevPush push;
    constexpr int kMode = (Class == PLTB_BSDF_ROUGHGRATING && !Rgb) ? kNeeFilter : kNeeFull;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t slot = runs.slot(e);
        if (st.rec_entry[slot] == kNoRecord) continue;
        stage_nee_post<Class, Rgb, kMode>(sc, st, slot, true, push, GratingInsideDirect());
    }
}

__global__ void __launch_bounds__(128) k_nee_scan(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st) {
    const uint32_t n = st.counts[kCntScan];
    DevPush push;
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_nee_post<PLTB_BSDF_ROUGHGRATING, false, kNeeScan>(sc, st, st.shadow_q[e], true, push, GratingInsideDirect());
}
__global__ void __launch_bounds__(128) k_nee_eval(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st) {
    const uint32_t n = st.counts[kCntEval];
    if (blockIdx.x == 0 && threadIdx.x == 0 && n) atomicAdd(st.counters + 6, (unsigned long long) n);
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_nee_eval(sc, st, st.queue_eval[e]);
}

// plt_solve_phase accumulation (plt.py:191-216), one thread per sample at full occupancy: the records of a path sit one DRAM
// round trip apart (depth-major arrays), so the loads are issued four vertices at a time before the ordered sum consumes them.
// The sum L is left in L_out[0] (PassState) for the splat kernels; optional per-sample dump.
__global__ void __launch_bounds__(256) k_gather(const __grid_constant__ PassState st, const __grid_constant__ PassParams pp,
                                                float *dump_wl, float *dump_L, uint32_t *dump_rays, uint64_t dump_cap) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= pp.count) return;
    const uint32_t n = st.n_vertices[local], m = st.v_mask[local];   // m: the terms that exist (the others are exact zeros)
    S4 L = s4(0.f);
    for (uint32_t base = 0; base < n; base += 4) {
        float4 em[4], ne[4];
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j)
            if (base + j < n) {
                const size_t r = (size_t) (base + j) * st.chunk + local;
                if (m >> (base + j) & 1u) em[j] = st.v_em[r];
                if (m >> (16 + base + j) & 1u) ne[j] = st.v_nee[r];
            }
#pragma unroll
        for (uint32_t j = 0; j < 4; ++j)
            if (base + j < n) {
                if (m >> (base + j) & 1u) L = L + f4_to_s4(em[j]);
                if (m >> (16 + base + j) & 1u) L = L + f4_to_s4(ne[j]);
            }
    }
    st.L_out[0][local] = s4_to_f4(L);
    if (dump_cap) {
        uint32_t pix, s_rel;
        split_work(pp, pp.first + local, pix, s_rel);
        const uint64_t di = (uint64_t) pix * pp.spp_count + s_rel;   // dump order: pixel-major, then sample (include/pltb.h)
        if (di < dump_cap) {
            if (dump_L) reinterpret_cast<float4 *>(dump_L)[di] = s4_to_f4(L);
            if (dump_wl) reinterpret_cast<float4 *>(dump_wl)[di] = st.wl[local];
            if (dump_rays) { dump_rays[2 * di] = n; dump_rays[2 * di + 1] = st.n_shadow[local]; }
        }
    }
}

// generic film splat (box filter, filters wider than 5x5): one thread per sample, one vector atomic per tap
__global__ void __launch_bounds__(256) k_splat(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                               const __grid_constant__ PassParams pp, float *accum) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= pp.count) return;
    DevFilm film{ accum };
    float values[4];
    splat_values(sc, st, local, f4_to_s4(st.L_out[0][local]), values);
    splat_taps(sc, pp, values, st.film_pos[local], film);
}

// Film splat through per-lane register tiles and per-warp shared-memory strips.  A warp owns 32 consecutive window pixels of
// one row and up to kSplatS sample indices of the pass.  All samples of a lane belong to the same pixel, so their
// (2 NR + 1)^2 filter taps land on the same footprint: they are summed in registers first (no memory traffic at all), the 32
// footprints are then merged into the warp's (32 + 2 NR) x (2 NR + 1) strip in shared memory with plain read-modify-writes
// -- at a fixed tap (ys, xs) the 32 lanes touch 32 distinct columns, so a __syncwarp() between taps is all the ordering that
// is needed (no shared atomics, which are CAS loops for float) -- and the strip is flushed with one red.global.add.v4.f32 per
// strip pixel: 180 instead of 6400 film atomics per 256 samples for the 5x5 Gaussian.  Lanes whose position does not floor
// to their own pixel (x + u rounding up to x + 1 in float32), or whose pixel sits on another row than lane 0's (window width
// not a multiple of 32), take the per-tap global path.
#ifndef PLTB_SPLAT_S
#define PLTB_SPLAT_S 32   // sample indices per warp (8: 49.2, 16: 46.1, 32: 45.0 ms of k_splat per 2^30 samples at 32 spp per pass)
#endif
constexpr int kSplatWarps = 4, kSplatS = PLTB_SPLAT_S;
// G < 0: the RGBW film.  G = 0..3: float4 group G of the 16-channel Stokes film (k_gather_pol left the four Stokes components of L
// in L_out[0..3]): four launches of this kernel replace one thread per sample issuing 100 vector atomics
// (k_splat_stokes: 61.6 of the 229 ms of a polarized cbox-grating render at 128 spp).
template <int G> __device__ __forceinline__ void stokes_group_values(const DeviceScene &sc, const PassState &st, uint32_t local, const S4 &wl,
                                                                     const S4 &sw, float values[4]) {
    const float4 *comp[4] = { st.L_out[0], st.L_out[1], st.L_out[2], st.L_out[3] };
    if (G == 0) { splat_values(sc, f4_to_s4(comp[0][local]), wl, sw, values); return; }
    // AOV i = spectrum_to_srgb(S_i / pdf_rgb_spectrum(wavelengths)) (stokes.cpp:108-121), RGB variants: S_i itself; flattened channels
    // 3 i + c sit at film channel 4 + 3 i + c, i.e. group G holds flattened channels 4 (G - 1) .. 4 (G - 1) + 3
    float aov[2][3];
    constexpr int i0 = (4 * (G - 1)) / 3;          // first Stokes component this group touches (G = 1: 0, 2: 1, 3: 2)
#pragma unroll
    for (int k = 0; k < 2; ++k) {
        const S4 c = f4_to_s4(comp[i0 + k][local]);
        if (sc.rgb) { aov[k][0] = c.v[0]; aov[k][1] = c.v[1]; aov[k][2] = c.v[2]; continue; }
        S4 v;
        for (int q = 0; q < 4; ++q) { const float pdf = pdf_rgb_spectrum(wl.v[q]); v.v[q] = c.v[q] * (pdf != 0.f ? rcp(pdf) : 0.f); }
        spectrum_to_srgb(sc.cie_srgb, v, wl, aov[k]);
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) { const int f = 4 * (G - 1) + q, i = f / 3, c = f - 3 * i; values[q] = aov[i - i0][c]; }
}
template <int NR, int G = -1>
__global__ void __launch_bounds__(32 * kSplatWarps) k_splat_strip(const __grid_constant__ DeviceScene sc,
                                                                  const __grid_constant__ PassState st,
                                                                  const __grid_constant__ PassParams pp, float *accum,
                                                                  uint32_t s_lo, uint32_t s_hi, uint32_t n_strips, uint32_t n_warps) {
    constexpr int kTaps = 2 * NR + 1, kCols = 32 + 2 * NR, kPlane = kTaps * kCols;
    __shared__ float4 smem[kSplatWarps * kPlane];
    const int warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    const uint32_t gw = blockIdx.x * kSplatWarps + warp;
    if (gw >= n_warps) return;                       // whole warps leave together; there is no block-level barrier below
    const uint32_t strip = gw % n_strips, s_first = s_lo + (gw / n_strips) * kSplatS;
    const uint32_t pix = strip * 32 + lane;
    const bool valid_pix = pix < pp.n_pix;
    const uint32_t row = pix / pp.crop_w;
    const int x = (int) (pp.crop_x + (pix - row * pp.crop_w)), y = (int) (pp.crop_y + row);
    const int x0 = __shfl_sync(0xffffffffu, x, 0), y0 = __shfl_sync(0xffffffffu, y, 0);
    const bool same_row = valid_pix && y == y0;      // then x == x0 + lane
    constexpr int kStride = G < 0 ? 1 : 4, kOff = G < 0 ? 0 : G;
    typename std::conditional<(G < 0), DevFilm, DevFilmGroup<(G < 0 ? 0 : G)>>::type film{ accum };
    float4 acc[kTaps][kTaps];
#pragma unroll
    for (int ys = 0; ys < kTaps; ++ys)
#pragma unroll
        for (int xs = 0; xs < kTaps; ++xs) acc[ys][xs] = make_float4(0.f, 0.f, 0.f, 0.f);
    bool dirty = false;
    // per-sample inputs are fetched one iteration ahead: the loop is a chain of dependent DRAM reads otherwise
    struct Fetch { bool in; uint32_t local; float4 L, wl, sw; float2 pos; };
    auto fetch = [&](uint32_t k) {
        Fetch f;
        const uint64_t w = (uint64_t) (s_first + k) * pp.n_pix + pix;
        f.in = k < (uint32_t) kSplatS && s_first + k <= s_hi && valid_pix && w >= pp.first && w < pp.first + pp.count;
        if (f.in) {
            const uint32_t local = (uint32_t) (w - pp.first);
            f.local = local;
            if (G < 0) f.L = st.L_out[0][local];
            f.wl = st.wl[local]; f.sw = st.sensor_w[local]; f.pos = st.film_pos[local];
        }
        return f;
    };
    Fetch nxt = fetch(0);
#pragma unroll 1
    for (uint32_t k = 0; k < (uint32_t) kSplatS && s_first + k <= s_hi; ++k) {
        const Fetch cur = nxt;
        nxt = fetch(k + 1);
        if (!cur.in) continue;
        float values[4];
        if (G < 0) splat_values(sc, f4_to_s4(cur.L), f4_to_s4(cur.wl), f4_to_s4(cur.sw), values);
        else stokes_group_values<(G < 0 ? 0 : G)>(sc, st, cur.local, f4_to_s4(cur.wl), f4_to_s4(cur.sw), values);
        const float2 pos = cur.pos;
        if (!(same_row && (int) floorf(pos.x) == x && (int) floorf(pos.y) == y)) {
            splat_taps(sc, pp, values, pos, film);
            continue;
        }
        dirty = true;
        float wx[kTaps], wy[kTaps];
        float relx = ((float) (x - NR) + .5f) - pos.x, rely = ((float) (y - NR) + .5f) - pos.y;
#pragma unroll
        for (int t = 0; t < kTaps; ++t) {
            wx[t] = rfilter_eval(sc.film, relx); wy[t] = rfilter_eval(sc.film, rely);
            relx += 1; rely += 1;
        }
#pragma unroll
        for (int ys = 0; ys < kTaps; ++ys)
#pragma unroll
            for (int xs = 0; xs < kTaps; ++xs) {
                const float wgt = wy[ys] * wx[xs];
                float4 &a = acc[ys][xs];
                a.x += values[0] * wgt; a.y += values[1] * wgt; a.z += values[2] * wgt; a.w += values[3] * wgt;
            }
    }
    if (!__any_sync(0xffffffffu, dirty)) return;
    // merge the 32 register footprints into the warp's strip
    float4 *tile = smem + warp * kPlane;
    for (int i = lane; i < kPlane; i += 32) tile[i] = make_float4(0.f, 0.f, 0.f, 0.f);
    __syncwarp();
#pragma unroll
    for (int ys = 0; ys < kTaps; ++ys)
#pragma unroll
        for (int xs = 0; xs < kTaps; ++xs) {
            if (dirty) {
                float4 *t = tile + ys * kCols + lane + xs, v = *t;
                const float4 a = acc[ys][xs];
                v.x += a.x; v.y += a.y; v.z += a.z; v.w += a.w;
                *t = v;
            }
            __syncwarp();
        }
    const int W = (int) pp.film_w, H = (int) pp.film_h;
    for (int i = lane; i < kPlane; i += 32) {
        const int r = i / kCols, c = i - r * kCols;
        const int xx = x0 - NR + c, yy = y0 - NR + r;
        if (xx < 0 || yy < 0 || xx >= W || yy >= H) continue;
        const float4 v = tile[i];
        if (v.x != 0.f || v.y != 0.f || v.z != 0.f || v.w != 0.f)   // NaN != 0: poisoned taps reach the film as before
            atomicAdd(reinterpret_cast<float4 *>(accum) + kStride * ((size_t) yy * W + xx) + kOff, v);
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

// ---- polarized variants (PLTB_RENDER_POLARIZED): same queues and launch order, Mueller arithmetic in the shading / NEE kernels ----
template <int Class>
__global__ void __launch_bounds__(128) k_shade_pol(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                   uint32_t bounce) {
    const uint32_t n = st.counts[cnt_shade(bounce, Class)];
    const uint32_t *list = st.shade_q[Class];
    uint32_t cnt[4] = { 0, 0, 0, 0 };
    DevPush push;
    uint32_t slot_base, live_base;
    block_slot_base(st, bounce, Class, slot_base, live_base);
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x)
        stage_shade_pol<Class>(sc, st, list[e], bounce, slot_base + e, live_base + e, push, cnt);
    flush_counters(cnt, st.counters);
}

template <int Class, bool Rgb>
__global__ void __launch_bounds__(128) k_nee_post_pol(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st) {
    __shared__ ClassRuns runs;
    const uint32_t n = block_class_runs(st, (uint32_t) Class, runs);
    for (uint32_t e = blockIdx.x * blockDim.x + threadIdx.x; e < n; e += gridDim.x * blockDim.x) {
        const uint32_t slot = runs.slot(e);
        if (st.rec_entry[slot] == kNoRecord) continue;
        stage_nee_post_pol<Class, Rgb>(sc, st, slot);
    }
}

// plt_solve_phase accumulation for the first column of L, then StokesIntegrator::sample's rotation to the sensor frame
// (stokes.cpp:97-106) when `stokes`.  The four components are left in L_out[0..3] (PassState).
__global__ void __launch_bounds__(256) k_gather_pol(const __grid_constant__ PassState st, const __grid_constant__ PassParams pp, int stokes,
                                                    float *dump_wl, float *dump_L, uint32_t *dump_rays, float *dump_stokes, uint64_t dump_cap) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= pp.count) return;
    const uint32_t n = st.n_vertices[local], m = st.v_mask[local];   // m: the terms that exist (the others are exact zeros)
    S4 L[4] = { s4(0.f), s4(0.f), s4(0.f), s4(0.f) };
    for (uint32_t i = 0; i < n; ++i) {
        const size_t r = (size_t) i * st.chunk + local;
        if (m >> i & 1u) for (int c = 0; c < 4; ++c) L[c] = L[c] + f4_to_s4(st.pv_em[(size_t) c * st.nv + r]);
        if (m >> (16 + i) & 1u) L[0] = L[0] + f4_to_s4(st.v_nee[r]);
    }
    if (stokes) {
        const float theta = st.stokes_theta[local];
        const F44 R = StokesFrames::rotator(sinf(2.f * theta), cosf(2.f * theta));
        for (int k = 0; k < 4; ++k) {
            C4 x; for (int c = 0; c < 4; ++c) x.v[c] = L[c].v[k];
            x = f44_apply(R, x);
            for (int c = 0; c < 4; ++c) L[c].v[k] = x.v[c];
        }
    }
    for (int c = 0; c < 4; ++c) st.L_out[c][local] = s4_to_f4(L[c]);
    if (dump_cap) {
        uint32_t pix, s_rel;
        split_work(pp, pp.first + local, pix, s_rel);
        const uint64_t di = (uint64_t) pix * pp.spp_count + s_rel;
        if (di < dump_cap) {
            if (dump_L) reinterpret_cast<float4 *>(dump_L)[di] = s4_to_f4(L[0]);
            if (dump_wl) reinterpret_cast<float4 *>(dump_wl)[di] = st.wl[local];
            if (dump_rays) { dump_rays[2 * di] = n; dump_rays[2 * di + 1] = st.n_shadow[local]; }
            if (dump_stokes) for (int c = 0; c < 4; ++c) reinterpret_cast<float4 *>(dump_stokes)[4 * di + c] = s4_to_f4(L[c]);
        }
    }
}

// film splat with the 12 Stokes AOVs (16 channels per pixel: R G B W S0.RGB .. S3.RGB): one thread per sample, four vector
// atomics per tap.  AOV i = spectrum_to_srgb(S_i / pdf_rgb_spectrum(wavelengths)) (stokes.cpp:108-121); the main channels are
// spectrum_to_srgb(ray_weight * S_0) (integrator.cpp:456-470).
__global__ void __launch_bounds__(256) k_splat_stokes(const __grid_constant__ DeviceScene sc, const __grid_constant__ PassState st,
                                                      const __grid_constant__ PassParams pp, float *accum) {
    const uint32_t local = blockIdx.x * blockDim.x + threadIdx.x;
    if (local >= pp.count) return;
    const S4 wl = f4_to_s4(st.wl[local]);
    const S4 comp[4] = { f4_to_s4(st.L_out[0][local]), f4_to_s4(st.L_out[1][local]), f4_to_s4(st.L_out[2][local]), f4_to_s4(st.L_out[3][local]) };
    float values[16];
    splat_values(sc, comp[0], wl, f4_to_s4(st.sensor_w[local]), values);
    for (int i = 0; i < 4; ++i) {
        if (sc.rgb) {   // rgb = spec.entry(i, 0) (stokes.cpp:114-115)
            for (int c = 0; c < 3; ++c) values[4 + 3 * i + c] = comp[i].v[c];
            continue;
        }
        S4 v;
        for (int k = 0; k < 4; ++k) { const float pdf = pdf_rgb_spectrum(wl.v[k]); v.v[k] = comp[i].v[k] * (pdf != 0.f ? rcp(pdf) : 0.f); }
        spectrum_to_srgb(sc.cie_srgb, v, wl, values + 4 + 3 * i);
    }
    struct Film16 {
        float *accum;
        __device__ __forceinline__ void add(size_t pixel, const float *v, float w) {
            float4 *dst = reinterpret_cast<float4 *>(accum) + 4 * pixel;
#pragma unroll
            for (int q = 0; q < 4; ++q) atomicAdd(dst + q, make_float4(v[4 * q] * w, v[4 * q + 1] * w, v[4 * q + 2] * w, v[4 * q + 3] * w));
        }
    } film{ accum };
    splat_taps(sc, pp, values, st.film_pos[local], film);
}

// HDRFilm::develop with AOV channels: every channel but W divided by (w == 0 ? 1 : w)
__global__ void k_develop_n(const float *accum, float *out, uint64_t n_pixels, int nch) {
    uint64_t i = (uint64_t) blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n_pixels) return;
    const float w = accum[nch * i + 3], d = w == 0.f ? 1.f : w;
    int o = 0;
    for (int c = 0; c < nch; ++c) if (c != 3) out[(nch - 1) * i + o++] = accum[nch * i + c] / d;
}

thread_local LaunchHook g_launch_hook;
#define PLTB_AFTER(k) do { if (g_launch_hook.fn) g_launch_hook.fn((k), g_launch_hook.user); } while (0)

void launch_raygen(const DeviceScene &sc, const PassState &st, const PassParams &pp, cudaStream_t s) {
    k_raygen<<<(pp.count + PLTB_RAYGEN_THREADS - 1) / PLTB_RAYGEN_THREADS, PLTB_RAYGEN_THREADS, 0, s>>>(sc, st, pp); PLTB_AFTER(PLTB_K_RAYGEN);
}
void launch_trace(const DeviceScene &sc, const PassState &st, uint32_t bounce, int grid, cudaStream_t s) {
    k_trace<false><<<grid, 128, 0, s>>>(sc, st, bounce); PLTB_AFTER(PLTB_K_TRACE);
    k_classify<<<grid, 256, 0, s>>>(sc, st, bounce); PLTB_AFTER(PLTB_K_CLASSIFY);
}
// one instantiation per material class (BSDF plugin type); `cls` selects it at run time
template <template <int> class Launch, class... Args> static void dispatch_class(int cls, Args &&...args) {
    switch (cls) {
        case PLTB_BSDF_NULL: Launch<PLTB_BSDF_NULL>::run(args...); break;
        case PLTB_BSDF_DIFFUSE: Launch<PLTB_BSDF_DIFFUSE>::run(args...); break;
        case PLTB_BSDF_CONDUCTOR: Launch<PLTB_BSDF_CONDUCTOR>::run(args...); break;
        case PLTB_BSDF_DIELECTRIC: Launch<PLTB_BSDF_DIELECTRIC>::run(args...); break;
        case PLTB_BSDF_ROUGHCONDUCTOR: Launch<PLTB_BSDF_ROUGHCONDUCTOR>::run(args...); break;
        case PLTB_BSDF_ROUGHGRATING: Launch<PLTB_BSDF_ROUGHGRATING>::run(args...); break;
        case PLTB_BSDF_PRINCIPLED: Launch<PLTB_BSDF_PRINCIPLED>::run(args...); break;
        default: break;
    }
}
template <int C> struct LaunchShade {
    static void run(const DeviceScene &sc, const PassState &st, uint32_t bounce, int grid, cudaStream_t s) {
        if (st.pv_w) k_shade_pol<C><<<grid, 128, 0, s>>>(sc, st, bounce);   // polarized variant
        else k_shade<C><<<grid, 128, 0, s>>>(sc, st, bounce);
    }
};
template <int C> struct LaunchNeePost {
    static void run(const DeviceScene &sc, const PassState &st, int grid, cudaStream_t s) {
        // the RGB variants differ in roughgrating's evaluation only (its RGB fallback): one more instantiation of that class
        constexpr bool kGrating = C == PLTB_BSDF_ROUGHGRATING;
        if (kGrating && sc.rgb) {
            if (st.pv_w) k_nee_post_pol<C, kGrating><<<grid, 128, 0, s>>>(sc, st);
            else k_nee_post<C, kGrating><<<grid, 128, 0, s>>>(sc, st);
        } else if (st.pv_w) k_nee_post_pol<C, false><<<grid, 128, 0, s>>>(sc, st);   // polarized: grating scan and evaluation in one kernel
        else k_nee_post<C, false><<<grid, 128, 0, s>>>(sc, st);
    }
};
void launch_shade(const DeviceScene &sc, const PassState &st, uint32_t bounce, int cls, int grid, cudaStream_t s) {
    dispatch_class<LaunchShade>(cls, sc, st, bounce, grid, s);
    PLTB_AFTER(cls == PLTB_BSDF_ROUGHGRATING ? PLTB_K_SHADE_GRATING : PLTB_K_SHADE_OTHER);
}
// emitter sampling for both queues, then all shadow rays in one persistent-warp launch
void launch_nee_light(const DeviceScene &sc, const PassState &st, uint32_t class_mask, int grid, cudaStream_t s) {
    for (int cls = kClasses - 1; cls >= 0; --cls)
        if (class_mask >> cls & 1u) { k_nee_light<<<grid, 256, 0, s>>>(sc, st, cls); PLTB_AFTER(PLTB_K_NEE_LIGHT); }
    k_trace<true><<<grid, 128, 0, s>>>(sc, st, 0); PLTB_AFTER(PLTB_K_TRACE_SHADOW);
}
void launch_nee_post(const DeviceScene &sc, const PassState &st, int cls, int grid, cudaStream_t s) {
    dispatch_class<LaunchNeePost>(cls, sc, st, grid, s);
    PLTB_AFTER(cls == PLTB_BSDF_ROUGHGRATING ? PLTB_K_NEE_POST_GRATING : PLTB_K_NEE_POST_OTHER);
    if (cls == PLTB_BSDF_ROUGHGRATING && !st.pv_w && !sc.rgb) {
        k_nee_scan<<<grid, 128, 0, s>>>(sc, st); PLTB_AFTER(PLTB_K_NEE_POST_GRATING);
        k_nee_eval<<<grid, 128, 0, s>>>(sc, st); PLTB_AFTER(PLTB_K_NEE_EVAL);
    }
}
void launch_splat(const DeviceScene &sc, const PassState &st, const PassParams &pp, float *accum, int do_splat, float *dump_wl,
                  float *dump_L, uint32_t *dump_rays, float *dump_stokes, int stokes, uint64_t dump_cap, cudaStream_t s) {
    if (st.pv_w)
        k_gather_pol<<<(pp.count + 255) / 256, 256, 0, s>>>(st, pp, stokes, dump_wl, dump_L, st.n_shadow ? dump_rays : nullptr, dump_stokes,
                                                            (dump_wl || dump_L || dump_rays || dump_stokes) ? dump_cap : 0);
    else
        k_gather<<<(pp.count + 255) / 256, 256, 0, s>>>(st, pp, dump_wl, dump_L, st.n_shadow ? dump_rays : nullptr,
                                                        (dump_wl || dump_L || dump_rays) ? dump_cap : 0);
    PLTB_AFTER(PLTB_K_GATHER);
    if (!do_splat) return;
    // separable filters with 3x3 (tent, r = 1) or 5x5 (gaussian, r = 2) taps go through the register tiles / shared strips
    const int nr = sc.film.rfilter == PLTB_RFILTER_BOX ? 0 : (int) ceilf(sc.film.radius - .5f);
    if ((nr == 1 || nr == 2) && !getenv("PLTB_SPLAT_DIRECT")) {
        const uint32_t s_lo = (uint32_t) (pp.first / pp.n_pix), s_hi = (uint32_t) ((pp.first + pp.count - 1) / pp.n_pix);
        const uint32_t n_strips = (pp.n_pix + 31) / 32, n_chunks = (s_hi - s_lo) / kSplatS + 1;
        const uint32_t n_warps = n_strips * n_chunks, grid = (n_warps + kSplatWarps - 1) / kSplatWarps;
#define PLTB_STRIP(NR, G) k_splat_strip<NR, G><<<grid, 32 * kSplatWarps, 0, s>>>(sc, st, pp, accum, s_lo, s_hi, n_strips, n_warps)
        if (stokes) {   // one launch per float4 group of the 16-channel film
            if (nr == 1) { PLTB_STRIP(1, 0); PLTB_STRIP(1, 1); PLTB_STRIP(1, 2); PLTB_STRIP(1, 3); }
            else { PLTB_STRIP(2, 0); PLTB_STRIP(2, 1); PLTB_STRIP(2, 2); PLTB_STRIP(2, 3); }
        } else if (nr == 1) PLTB_STRIP(1, -1);
        else PLTB_STRIP(2, -1);
#undef PLTB_STRIP
        PLTB_AFTER(PLTB_K_SPLAT);
        return;
    }
    if (stokes) {
        k_splat_stokes<<<(pp.count + 255) / 256, 256, 0, s>>>(sc, st, pp, accum); PLTB_AFTER(PLTB_K_SPLAT);
        return;
    }
    k_splat<<<(pp.count + 255) / 256, 256, 0, s>>>(sc, st, pp, accum); PLTB_AFTER(PLTB_K_SPLAT);
}
void launch_develop(const float *accum, float *rgb, uint64_t n_pixels, int nch, cudaStream_t s) {
    if (nch == 4) k_develop<<<(unsigned) ((n_pixels + 255) / 256), 256, 0, s>>>(reinterpret_cast<const float4 *>(accum), rgb, n_pixels);
    else k_develop_n<<<(unsigned) ((n_pixels + 255) / 256), 256, 0, s>>>(accum, rgb, n_pixels, nch);
}

}  // namespace pltb
