// Probe: what would ONE local-training step cost if its three batch GEMMs ran on tcgen05 instead of mma.sync?
//
// The training kernel (csrc/train_kernel.cu) runs layer 1, the output layer and the data gradient of a 12-row batch
// as error-compensated 3xTF32 on warp-level mma.sync (measured 1.1-1.4 k cycles per phase).  The tcgen05 form that
// was designed on paper puts the WEIGHTS on the M side (M = 128 lanes of TMEM) and the batch on the N side (N = 16):
//     layer 1        D[128 x 16] = W1 (27 -> 128 rows, K = 128) . X^T          16 k-steps x 3 products = 48 MMAs
//     output layer   D[128 x 16] = W4 (115 -> 128 rows, K = 32) . h2^T          4 k-steps x 3           = 12 MMAs
//     data gradient  D[128 x 16] = W4^T (27 -> 128 rows, K = 128) . dX^T       16 k-steps x 3           = 48 MMAs
// Every step also has to (a) write the batch-side operand (X, h2, dX: 16 x K fp32) as TF32 head + remainder images in
// the UMMA shared-memory layout, (b) read D back with tcgen05.ld (one lane per weight row), and (c) re-split the three
// weight images after the Adam update (they change every step).  This probe times those pieces on one CTA with the
// same instruction sequences the kernel would use (values are dummies):
//   [0] 48 MMAs issue -> retire     [1] + tcgen05.ld of D (4 warps)        (layer 1 / data gradient shape)
//   [2] 12 MMAs issue -> retire     [3] + tcgen05.ld of D                  (output layer shape)
//   [4] batch-side operand image: 16 x 128 floats -> hi/lo swizzled + fence.proxy.async (128 threads)
//   [5] weight-side images: 27x115 + 115x27 + 27x115(T) floats -> hi/lo swizzled + fence (512 threads' worth, 128 here x4)
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/train_step_tc_probe.cu -o build/train_step_tc_probe -lcuda
#include <cstdio>
#include "tc_common.cuh"

using namespace fm;
using namespace fm::tc;

struct PSmem {
    alignas(1024) unsigned char a[2][4 * XBLK];      // weight image hi | lo: 128 rows x 128 floats
    alignas(1024) unsigned char b[2][4 * 16 * 128];  // batch image hi | lo: 16 rows x 128 floats (4 K-blocks of 16 x 128 B)
    alignas(8) uint64_t bar;
    uint32_t tmem_base;
    float src[128 * 128];                             // fp32 source the images are split from
};

template <int KSTEPS>
__device__ __forceinline__ void issue_gemm(const PSmem& S, uint32_t tmem, uint64_t da, uint64_t dalo, uint64_t db, uint64_t dblo) {
    constexpr uint32_t idesc = idesc_tf32(128, 16);
#pragma unroll
    for (int kk = 0; kk < KSTEPS; ++kk) {
        const int kb = kk >> 2, ks = kk & 3;
        const uint64_t oa = (uint64_t)((kb * XBLK + ks * 32) >> 4), ob = (uint64_t)((kb * 16 * 128 + ks * 32) >> 4);
        mma_ss(tmem, da + oa, db + ob, idesc, kk ? 1u : 0u);
        mma_ss(tmem, da + oa, dblo + ob, idesc, 1u);
        mma_ss(tmem, dalo + oa, db + ob, idesc, 1u);
    }
}

__global__ void __launch_bounds__(160, 1) probe(long long* out) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    PSmem& S = *reinterpret_cast<PSmem*>(smem_dyn);
    const int tid = threadIdx.x, warp = tid >> 5, lane = tid & 31;
    for (int i = tid; i < (int)sizeof(PSmem) / 4; i += 160) reinterpret_cast<uint32_t*>(smem_dyn)[i] = 0u;
    __syncthreads();
    if (tid == 0) { mbar_init(&S.bar, 1); fence_mbar_init(); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 4) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(32));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    long long acc[6] = {0, 0, 0, 0, 0, 0};
    uint32_t phase = 0;
    constexpr int REPS = 8;
    for (int rep = 0; rep < REPS; ++rep) {
        for (int shape = 0; shape < 2; ++shape) {
            __syncthreads();
            const long long t0 = clock64();
            if (warp == 4) {
                if (elect_one()) {
                    const uint64_t da = smem_desc_sw128(S.a[0]), dalo = smem_desc_sw128(S.a[1]);
                    const uint64_t db = smem_desc_sw128(S.b[0]), dblo = smem_desc_sw128(S.b[1]);
                    if (shape == 0) issue_gemm<16>(S, tmem, da, dalo, db, dblo); else issue_gemm<4>(S, tmem, da, dalo, db, dblo);
                    tc_commit(&S.bar);
                }
                __syncwarp();
            }
            if (warp < 4) {
                mbar_wait(&S.bar, phase);
                tc_fence_after();
                const long long t1 = clock64();
                uint32_t v[16];
                asm volatile("tcgen05.ld.sync.aligned.32x32b.x16.b32 {%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];\n\t"
                             "tcgen05.wait::ld.sync.aligned;\n"
                             : "=r"(v[0]), "=r"(v[1]), "=r"(v[2]), "=r"(v[3]), "=r"(v[4]), "=r"(v[5]), "=r"(v[6]), "=r"(v[7]), "=r"(v[8]),
                               "=r"(v[9]), "=r"(v[10]), "=r"(v[11]), "=r"(v[12]), "=r"(v[13]), "=r"(v[14]), "=r"(v[15])
                             : "r"(tmem + ((uint32_t)(warp * 32) << 16)) : "memory");
                uint32_t x = 0;
                for (int j = 0; j < 16; ++j) x ^= v[j];
                if (x == 0xdeadbeefu) S.src[tid] = 1.f;          // keep the loads alive
                const long long t2 = clock64();
                if (tid == 0) { acc[shape * 2] += t1 - t0; acc[shape * 2 + 1] += t2 - t0; }
            }
            phase ^= 1u;
        }
        // ---- batch-side operand image: 16 x 128 floats split into hi / lo, swizzled K-major ----
        __syncthreads();
        {
            const long long t0 = clock64();
            if (tid < 128) {
                for (int e = tid; e < 16 * 128; e += 128) {
                    const int n = e >> 7, k = e & 127;
                    const float x = S.src[e];
                    const uint32_t hi = trunc_tf32(x);
                    const int off = (k >> 5) * (16 * 128) + sw128(n, k & 31);
                    *reinterpret_cast<uint32_t*>(S.b[0] + off) = hi;
                    *reinterpret_cast<float*>(S.b[1] + off) = x - __uint_as_float(hi);
                }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            __syncthreads();
            if (tid == 0) acc[4] += clock64() - t0;
        }
        // ---- weight-side images after an Adam step: 3 x (27 x 115) floats, 128 threads (the kernel has 512: / 4) ----
        {
            const long long t0 = clock64();
            if (tid < 128) {
                for (int img = 0; img < 3; ++img)
                    for (int e = tid; e < 27 * 115; e += 128) {
                        const int n = img == 1 ? e / 27 : e / 115, k = img == 1 ? e % 27 : e % 115;
                        const float x = S.src[(n * 128 + k) & 16383];
                        const uint32_t hi = trunc_tf32(x);
                        const int off = (k >> 5) * XBLK + sw128(n, k & 31);
                        *reinterpret_cast<uint32_t*>(S.a[0] + off) = hi;
                        *reinterpret_cast<float*>(S.a[1] + off) = x - __uint_as_float(hi);
                    }
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
            }
            __syncthreads();
            if (tid == 0) acc[5] += clock64() - t0;
        }
    }
    if (tid == 0) for (int i = 0; i < 6; ++i) out[i] = acc[i] / REPS;
    tc_fence_before();
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(32));
}

int main() {
    long long* out;
    cudaMalloc(&out, 64);
    cudaFuncSetAttribute(probe, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PSmem));
    for (int it = 0; it < 2; ++it) {
        probe<<<1, 160, sizeof(PSmem)>>>(out);
        long long h[6];
        if (cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(cudaGetLastError())); return 1; }
        if (it == 1) {
            printf("tcgen05 training-step probe (cycles, mean of 8 repetitions, one CTA)\n");
            printf("  K=128 GEMM (48 MMAs, M128 N16 K8 3xTF32): issue->retire %lld, + tcgen05.ld %lld   (layer 1; data gradient)\n", h[0], h[1]);
            printf("  K=32  GEMM (12 MMAs):                      issue->retire %lld, + tcgen05.ld %lld   (output layer)\n", h[2], h[3]);
            printf("  batch-side operand image (16x128 hi/lo, 128 threads): %lld\n", h[4]);
            printf("  weight-side images (3 x 27x115 hi/lo, 128 threads):   %lld  (~ /4 with the kernel's 512 threads)\n", h[5]);
            const long long step = 2 * (h[1] + h[4]) + (h[3] + h[4]) + h[5] / 4;
            printf("  => three GEMM phases of one step on tcgen05: ~%lld cycles  (mma.sync today: 1.2 k + 1.4 k + 1.1 k = 3.7 k)\n", step);
        }
    }
    return 0;
}
