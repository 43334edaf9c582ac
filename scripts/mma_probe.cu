// Micro-probe: cost of back-to-back tcgen05.mma (kind::tf32, M = 128, K = 8) as a function of N, of the
// A-operand source (TMEM = TS, shared memory = SS) and of how many independent accumulators the chain
// alternates over.  Answers "is a small-N MMA bound by math, by the A read or by the accumulate dependency?"
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/mma_probe.cu -o build/mma_probe -lcuda
#include <cstdio>
#include "tc_common.cuh"

using namespace fm;
using namespace fm::tc;

struct ProbeSmem {
    alignas(1024) unsigned char a[4 * XBLK];       // 128 rows x 128 floats, 4 K-blocks
    alignas(1024) unsigned char b[4 * 256 * 128];  // up to N = 256 rows x 4 K-blocks
    alignas(8) uint64_t bar;
    uint32_t tmem_base;
};

// MODE: 0 = TS, 1 = SS.  N_ACC accumulators, each N columns wide (N_ACC * N <= 256), A in TMEM columns 256..383.
// Everything about the issue sequence is a compile-time constant and the issuing thread is chosen with elect.sync,
// so the UTCHMMA instructions are emitted back to back (behind `tid == 0` ptxas wraps every MMA in an
// election loop and the probe measures the loop, ~150 cycles per MMA, instead of the tensor pipe).
template <int MODE, int N, int N_ACC>
__global__ void __launch_bounds__(128, 1) probe(long long* out) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    ProbeSmem& S = *reinterpret_cast<ProbeSmem*>(smem_dyn);
    const int tid = threadIdx.x, warp = tid >> 5;
    for (int i = tid; i < (int)(sizeof(S.a) + sizeof(S.b)) / 4; i += 128) reinterpret_cast<uint32_t*>(smem_dyn)[i] = 0u;
    if (tid == 0) { mbar_init(&S.bar, 1); fence_mbar_init(); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 0) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    {   // zero the A columns (lane quarter of this warp)
        uint32_t v[32];
        for (int j = 0; j < 32; ++j) v[j] = 0u;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        for (int cb = 0; cb < 4; ++cb) tmem_st32(lane_base + 256 + cb * 32, v);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    constexpr int N_MMA = 96;
    if (warp == 0 && elect_one()) {
        constexpr uint32_t idesc = idesc_tf32(128, N);
        const uint64_t da = smem_desc_sw128(S.a), db = smem_desc_sw128(S.b);
        for (int rep = 0; rep < 3; ++rep) {
            const long long t0 = clock64();
#pragma unroll
            for (int i = 0; i < N_MMA; ++i) {
                const int kk = i & 15, kb = kk >> 2, ks = kk & 3;
                const uint32_t d = tmem + (uint32_t)((i % N_ACC) * N);
                const uint64_t bd = db + (uint64_t)((kb * N * 128 + ks * 32) >> 4);
                if (MODE == 0) mma_ts(d, tmem + 256 + kk * 8, bd, idesc, i >= N_ACC ? 1u : 0u);
                else mma_ss(d, da + (uint64_t)((kb * XBLK + ks * 32) >> 4), bd, idesc, i >= N_ACC ? 1u : 0u);
            }
            const long long t1 = clock64();
            tc_commit(&S.bar);
            mbar_wait(&S.bar, (uint32_t)(rep & 1));
            const long long t2 = clock64();
            out[rep * 2] = t1 - t0;
            out[rep * 2 + 1] = t2 - t0;
        }
    }
    __syncwarp();
    tc_fence_before();
    __syncthreads();
    if (warp == 0) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

template <int MODE, int N, int N_ACC>
void run(long long* out) {
    cudaFuncSetAttribute(probe<MODE, N, N_ACC>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(ProbeSmem));
    probe<MODE, N, N_ACC><<<1, 128, sizeof(ProbeSmem)>>>(out);
    long long h[6];
    if (cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost) != cudaSuccess) { printf("CUDA error: %s\n", cudaGetErrorString(cudaGetLastError())); return; }
    printf("%s  %3d  %d  %7.1f  %7.1f\n", MODE ? "SS" : "TS", N, N_ACC, (double)h[4] / 96, (double)h[5] / 96);
}

int main() {
    long long* out;
    cudaMalloc(&out, 64);
    printf("mode  N  n_acc  cycles/MMA (issue loop)  cycles/MMA (to completion)   [kind::tf32, M = 128, K = 8]\n");
    run<0, 32, 1>(out); run<0, 32, 2>(out); run<0, 32, 4>(out); run<0, 64, 1>(out); run<0, 64, 2>(out);
    run<0, 128, 1>(out); run<0, 128, 2>(out); run<0, 256, 1>(out);
    run<1, 32, 1>(out); run<1, 32, 4>(out); run<1, 64, 1>(out); run<1, 128, 1>(out); run<1, 256, 1>(out);
    return 0;
}
