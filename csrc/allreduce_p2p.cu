// Weighted model reduction across GPUs for WIDE models (BASELINE.json config 5: widened SAE, 1-512 MB of
// parameters per client).  For the 27 KB model the merge is latency bound and merge_kernel.cu simply pulls
// every selected row from its owner.  When a model is megabytes, bytes over NVLink decide, so the reduction
// is done bandwidth-optimally, entirely with in-kernel peer loads (no NCCL, no staging copies):
//
//   phase 1  partial_r   = sum over local models k of w_k * theta_k            (local HBM reads)
//   phase 2  reduced_r[slice_r] = sum over ranks q of partial_q[slice_r]        (P2P loads: (N-1)/N * P in)
//   phase 3  out[slice_q] = reduced_q[slice_q] for every q                      (P2P loads: (N-1)/N * P in)
//
// i.e. a reduce-scatter and an all-gather by pulling, 2 (N-1)/N * P bytes per GPU instead of (N-1) * P for
// the naive pull-everything (also provided: mode 0).  Phases are separate launches on each rank's stream;
// cross-rank ordering uses the control-block arrival counters (release reduction / acquire spin) of
// merge_kernel.cu, so a phase starts reading a peer's buffer only after that peer has finished writing it.
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

__device__ __forceinline__ void ar_signal(unsigned* addr) {
    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(addr) : "memory");
}
__device__ __forceinline__ unsigned ar_load(const unsigned* addr) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
__device__ __forceinline__ void ar_wait(const unsigned* arrive, unsigned expect) {
    if (arrive == nullptr) return;
    if (threadIdx.x == 0) while ((int)(ar_load(arrive) - expect) < 0) __nanosleep(64);
    __syncthreads();
}
// last block of the grid bumps `which` on every rank's control block
__device__ __forceinline__ void ar_signal_all(float* const* peer_ctl, int world, int which, unsigned* grid_counter) {
    if (peer_ctl == nullptr) return;
    __shared__ unsigned s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0) s_last = atomicAdd(grid_counter, 1u);
    __syncthreads();
    if (s_last == gridDim.x - 1) {
        if (threadIdx.x == 0) *grid_counter = 0u;
        if ((int)threadIdx.x < world)
            ar_signal(reinterpret_cast<unsigned*>(peer_ctl[threadIdx.x] + 2 * FM_MAX_SEL) + which);
    }
}

// out[p] = sum_k w[k] * rows[k][p], rows given as a device table of pointers (local or peer-mapped)
__global__ void __launch_bounds__(256) weighted_rows_kernel(const float* const* __restrict__ rows, const float* __restrict__ w,
                                                            int n, float* __restrict__ out, size_t p4, const unsigned* arrive,
                                                            unsigned expect, float* const* peer_ctl, int world, int which,
                                                            unsigned* grid_counter) {
    ar_wait(arrive, expect);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < p4; i += (size_t)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        int k = 0;
        for (; k + 4 <= n; k += 4) {
            float4 v[4];
#pragma unroll
            for (int u = 0; u < 4; ++u) v[u] = reinterpret_cast<const float4*>(rows[k + u])[i];
#pragma unroll
            for (int u = 0; u < 4; ++u) {
                const float wk = w[k + u];
                acc.x = fmaf(wk, v[u].x, acc.x); acc.y = fmaf(wk, v[u].y, acc.y);
                acc.z = fmaf(wk, v[u].z, acc.z); acc.w = fmaf(wk, v[u].w, acc.w);
            }
        }
        for (; k < n; ++k) {
            const float4 v = reinterpret_cast<const float4*>(rows[k])[i];
            const float wk = w[k];
            acc.x = fmaf(wk, v.x, acc.x); acc.y = fmaf(wk, v.y, acc.y); acc.z = fmaf(wk, v.z, acc.z); acc.w = fmaf(wk, v.w, acc.w);
        }
        reinterpret_cast<float4*>(out)[i] = acc;
    }
    ar_signal_all(peer_ctl, world, which, grid_counter);
}

// phase 2: reduced[slice] = sum_q partial_q[slice]   (partials: peer-mapped, one per rank)
__global__ void __launch_bounds__(256) slice_reduce_kernel(const float* const* __restrict__ partials, int world, int rank,
                                                           float* __restrict__ reduced, size_t slice0_4, size_t slice_n4,
                                                           const unsigned* arrive, unsigned expect, float* const* peer_ctl,
                                                           int which, unsigned* grid_counter) {
    ar_wait(arrive, expect);
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < slice_n4; i += (size_t)gridDim.x * blockDim.x) {
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
#pragma unroll 4
        for (int q = 0; q < world; ++q) {                   // fixed rank order: every rank gets identical bits
            const float4 v = reinterpret_cast<const float4*>(partials[q])[slice0_4 + i];
            acc.x += v.x; acc.y += v.y; acc.z += v.z; acc.w += v.w;
        }
        reinterpret_cast<float4*>(reduced)[slice0_4 + i] = acc;
    }
    ar_signal_all(peer_ctl, world, which, grid_counter);
}

// phase 3: out[slice_q] = reduced_q[slice_q] for every rank q
__global__ void __launch_bounds__(256) slice_gather_kernel(const float* const* __restrict__ reduced, int world,
                                                           float* __restrict__ out, size_t p4, const unsigned* arrive,
                                                           unsigned expect) {
    ar_wait(arrive, expect);
    const size_t per = (p4 + world - 1) / world;
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < p4; i += (size_t)gridDim.x * blockDim.x) {
        const int q = (int)(i / per);
        reinterpret_cast<float4*>(out)[i] = reinterpret_cast<const float4*>(reduced[q])[i];
    }
}

}  // namespace fm

extern "C" int fm_weighted_rows(const float* const* rows, const float* w, int32_t n, float* out, int64_t n_floats,
                                const uint32_t* arrive, uint32_t expect, float* const* peer_ctl, int32_t world,
                                int32_t which, uint32_t* grid_counter, void* stream) {
    if (n_floats % 4) return (int)cudaErrorInvalidValue;
    const size_t p4 = (size_t)n_floats / 4;
    const int grid = (int)((p4 + 255) / 256 < 148 * 8 ? (p4 + 255) / 256 : 148 * 8);
    fm::weighted_rows_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(rows, w, n, out, p4, arrive, expect, peer_ctl, world,
                                                                   which, grid_counter);
    return (int)cudaGetLastError();
}

extern "C" int fm_slice_reduce(const float* const* partials, int32_t world, int32_t rank, float* reduced, int64_t n_floats,
                               const uint32_t* arrive, uint32_t expect, float* const* peer_ctl, int32_t which,
                               uint32_t* grid_counter, void* stream) {
    if (n_floats % 4) return (int)cudaErrorInvalidValue;
    const size_t p4 = (size_t)n_floats / 4, per = (p4 + world - 1) / world;
    const size_t s0 = per * rank, sn = s0 >= p4 ? 0 : (p4 - s0 < per ? p4 - s0 : per);
    const int grid = (int)((sn + 255) / 256 < 148 * 8 ? (sn + 255) / 256 : 148 * 8);
    fm::slice_reduce_kernel<<<grid > 0 ? grid : 1, 256, 0, (cudaStream_t)stream>>>(partials, world, rank, reduced, s0, sn, arrive,
                                                                              expect, peer_ctl, which, grid_counter);
    return (int)cudaGetLastError();
}

extern "C" int fm_slice_gather(const float* const* reduced, int32_t world, float* out, int64_t n_floats,
                               const uint32_t* arrive, uint32_t expect, void* stream) {
    if (n_floats % 4) return (int)cudaErrorInvalidValue;
    const size_t p4 = (size_t)n_floats / 4;
    const int grid = (int)((p4 + 255) / 256 < 148 * 8 ? (p4 + 255) / 256 : 148 * 8);
    fm::slice_gather_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reduced, world, out, p4, arrive, expect);
    return (int)cudaGetLastError();
}
