// Shared device helpers for the fedmse_b200 kernels (sm_100a only).
//
// Model: D -> H -> ReLU -> Z -> H -> ReLU -> D autoencoder (reference:
// src/Model/Shrink_Autoencoder.py:40-42, :95-97).  One client's parameters are a
// packed fp32 row of P_PAD floats in state_dict order (see layout.py).
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace fm {

constexpr int D = 115;   // features
constexpr int H = 27;    // hidden
constexpr int Z = 7;     // latent
constexpr int XP = 128;  // row pitch (floats) of every data shard

// packed-row offsets (state_dict order)
constexpr int OFF_W1 = 0;                 // [H][D]
constexpr int OFF_B1 = OFF_W1 + H * D;    // [H]
constexpr int OFF_W2 = OFF_B1 + H;        // [Z][H]
constexpr int OFF_B2 = OFF_W2 + Z * H;    // [Z]
constexpr int OFF_W3 = OFF_B2 + Z;        // [H][Z]
constexpr int OFF_B3 = OFF_W3 + H * Z;    // [H]
constexpr int OFF_W4 = OFF_B3 + H;        // [D][H]
constexpr int OFF_B4 = OFF_W4 + D * H;    // [D]
constexpr int NP = OFF_B4 + D;            // 6764
constexpr int P_PAD = (NP + 31) / 32 * 32;  // 6784
static_assert(NP == 6764 && P_PAD == 6784, "layout drifted from layout.py");

// which of the 8 tensors a packed index belongs to
__host__ __device__ constexpr int tensor_of(int p) {
    return p < OFF_B1 ? 0 : p < OFF_W2 ? 1 : p < OFF_B2 ? 2 : p < OFF_W3 ? 3
         : p < OFF_B3 ? 4 : p < OFF_W4 ? 5 : p < OFF_B4 ? 6 : 7;
}

__device__ __forceinline__ float warp_sum(float v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}
__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// Block-wide sum; result valid in every thread. `scratch` needs >= 33 entries.
template <typename T>
__device__ __forceinline__ T block_sum(T v, T* scratch) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
    v = warp_sum(v);
    __syncthreads();
    if (lane == 0) scratch[w] = v;
    __syncthreads();
    if (w == 0) {
        T t = lane < nw ? scratch[lane] : T(0);
        t = warp_sum(t);
        if (lane == 0) scratch[32] = t;
    }
    __syncthreads();
    return scratch[32];
}

// ---- mbarrier + bulk async copy (TMA unit, 1-D form; SASS: UBLKCP / SYNCS) ----------------
__device__ __forceinline__ uint32_t smem_u32(const void* p) {
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count) {
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count));
}
__device__ __forceinline__ void fence_mbar_init() {
    asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
}
__device__ __forceinline__ void fence_proxy_async() {
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes) {
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity) {
    asm volatile(
        "{\n\t"
        ".reg .pred p;\n\t"
        "WAIT_%=:\n\t"
        "mbarrier.try_wait.parity.shared::cta.b64 p, [%0], %1;\n\t"
        "@p bra DONE_%=;\n\t"
        "bra WAIT_%=;\n\t"
        "DONE_%=:\n\t"
        "}" ::"r"(smem_u32(bar)),
        "r"(parity)
        : "memory");
}
// global -> shared bulk copy; bytes must be a multiple of 16, both addresses 16-B aligned.
__device__ __forceinline__ void bulk_g2s(void* dst_smem, const void* src_gmem, uint32_t bytes, uint64_t* bar) {
    asm volatile("cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];" ::"r"(
                     smem_u32(dst_smem)),
                 "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
                 : "memory");
}

}  // namespace fm
