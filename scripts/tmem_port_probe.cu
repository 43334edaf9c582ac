// Probe: does tcgen05.ld traffic (the epilogue warps reading accumulators out of tensor memory) slow the tensor pipe down?
//
// The fused scoring kernel reads 96 KB per (tile, model) pair out of TMEM (two ReLU hops of 16 KB, the squared-error warps'
// 64 KB) while the tensor pipe needs 717 cycles per pair; alone the MMA side sustains 786 cycles per pair, with the
// epilogues 1,237.  This probe runs, on one CTA,
//   (a) LD warps only: 4 or 8 warps (1 or 2 per SM sub-partition) issuing tcgen05.ld.32x32b.x32 back to back
//   (b) the MMA thread only: trains of N = 128 or N = 32 MMAs (A in TMEM)
//   (c) both at once
// and prints bytes / clk of the loads and cycles per MMA in each case.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/tmem_port_probe.cu -o build/tmem_port_probe
#include <cstdio>
#include <cstdlib>
#include "tc_common.cuh"

using namespace fm;
using namespace fm::tc;

__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_ts_f16(uint32_t d, uint32_t a, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
                 ::"r"(d), "r"(a), "l"(bdesc), "r"(idesc), "r"(accum), "r"(0u) : "memory");
}

__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }

struct PSmem {
    alignas(1024) unsigned char b[128 * 128];
    alignas(8) uint64_t bar;
    uint32_t tmem_base;
    int stop;
};

constexpr int R_MMA = 2048, R_LD = 1024;

// mode bit 0: LD warps run, bit 1: MMA thread runs; n_ld_warps = 4 or 8; N = MMA width
template <int N>
__global__ void __launch_bounds__(320, 1) probe(long long* out, int mode, int n_ld_warps) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    PSmem& S = *reinterpret_cast<PSmem*>(smem_dyn);
    const int tid = threadIdx.x, warp = __shfl_sync(0xffffffffu, tid >> 5, 0), lane = tid & 31;
    for (int i = tid; i < (int)sizeof(PSmem) / 4; i += 320) reinterpret_cast<uint32_t*>(smem_dyn)[i] = 0u;
    __syncthreads();
    if (tid == 0) { mbar_init(&S.bar, 1); fence_mbar_init(); }
    asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
    if (warp == 9) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(512));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    if (warp == 8) {
        if ((mode & 2) && elect_one()) {
            constexpr uint32_t idesc = idesc_f16(128, N);
            const uint64_t db = smem_desc_sw128(S.b);
            const long long t0 = clock64();
            if (mode & 8) {               // the kernel's layer-1 pattern: A walks over 2 x 64 columns (x hi | x lo), 3 MMAs per k-step
                for (int g = 0; g < R_MMA / 24; ++g) {
                    const uint32_t d = tmem + (uint32_t)((g & 1) ? 128 : 0);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        const uint64_t bh = db + (uint64_t)(((kk & 3) * 32) >> 4);
                        mma_ts_f16(d, tmem + 384 + kk * 8, bh, idesc, 1u);
                        mma_ts_f16(d, tmem + 384 + kk * 8, bh, idesc, 1u);
                        mma_ts_f16(d, tmem + 448 + kk * 8, bh, idesc, 1u);
                    }
                }
            } else {
#pragma unroll 8
                for (int i = 0; i < R_MMA; ++i)
                    mma_ts_f16(tmem + (uint32_t)((i & 8) ? 128 : 0), tmem + 480, db + (uint64_t)(((i & 3) * 32) >> 4), idesc, 1u);
            }
            tc_commit(&S.bar);
            mbar_wait(&S.bar, 0);
            out[0] = clock64() - t0;
        }
        __syncwarp();
    } else if (warp < n_ld_warps && (mode & 1)) {
        // the loads read the accumulator columns the MMAs write (256..383 are idle columns: second variant below)
        const uint32_t base = tmem + ((uint32_t)((warp & 3) * 32) << 16) + ((mode & 4) ? 256u : 0u) + (warp >= 4 ? 64u : 0u);
        uint32_t v[32], x = 0;
        const long long t0 = clock64();
        for (int i = 0; i < R_LD; ++i) {
            if (mode & 16) {              // the kernel's ReLU hop: load 32 columns, convert, store 16 columns in place, wait, fence
                tc_fence_after();
                tmem_ld32(base + (uint32_t)((i & 1) * 32), v);
                uint32_t hi[16];
#pragma unroll
                for (int j = 0; j < 16; ++j) hi[j] = v[2 * j] ^ (v[2 * j + 1] >> 1);
                tmem_st16(base + (uint32_t)((i & 1) * 32), hi);
                tmem_st_wait();
                tc_fence_before();
                x ^= hi[3];
            } else {
                tmem_ld32(base + (uint32_t)((i & 1) * 32), v);
#pragma unroll
                for (int j = 0; j < 32; ++j) x ^= v[j];
            }
        }
        const long long t1 = clock64();
        if (x == 0xdeadbeefu) S.stop = 1;
        if (lane == 0) out[1 + warp] = t1 - t0;
    }
    tc_fence_before();
    __syncthreads();
    if (warp == 9) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(512));
}

template <int N>
void run(long long* out, int mode, int n_ld, const char* what) {
    cudaFuncSetAttribute(probe<N>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(PSmem) + 1024);
    cudaMemset(out, 0, 16 * sizeof(long long));
    probe<N><<<1, 320, sizeof(PSmem) + 1024>>>(out, mode, n_ld);
    cudaError_t e = cudaDeviceSynchronize();
    if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); exit(1); }
    long long h[16];
    cudaMemcpy(h, out, sizeof(h), cudaMemcpyDeviceToHost);
    printf("  %-64s", what);
    if (mode & 2) printf(" MMA N=%-3d %6.1f cycles each (nominal %d)", N, (double)h[0] / ((mode & 8) ? (R_MMA / 24) * 24 : R_MMA), N / 2);
    if (mode & 1) {
        long long mx = 0;
        for (int w = 0; w < n_ld; ++w) mx = h[1 + w] > mx ? h[1 + w] : mx;
        printf("  loads: %6.1f cycles per 4 KB tcgen05.ld.x32 per warp, %6.1f B/clk per SM", (double)mx / R_LD,
               (double)n_ld * R_LD * 4096.0 / (double)mx);
    }
    printf("\n");
}

int main() {
    long long* out;
    cudaMalloc(&out, 16 * sizeof(long long));
    run<128>(out, 1, 4, "loads only, 4 warps (one per sub-partition)");
    run<128>(out, 1, 8, "loads only, 8 warps (two per sub-partition)");
    run<128>(out, 2, 4, "MMAs only, N = 128");
    run<32>(out, 2, 4, "MMAs only, N = 32");
    run<128>(out, 3, 4, "N = 128 MMAs + 4 load warps reading the accumulator columns");
    run<128>(out, 3, 8, "N = 128 MMAs + 8 load warps reading the accumulator columns");
    run<128>(out, 7, 8, "N = 128 MMAs + 8 load warps reading OTHER columns");
    run<32>(out, 3, 4, "N = 32 MMAs + 4 load warps");
    run<32>(out, 3, 8, "N = 32 MMAs + 8 load warps");
    run<32>(out, 7, 8, "N = 32 MMAs + 8 load warps reading OTHER columns");
    run<32>(out, 2 | 8, 4, "N = 32 MMAs, A walking over 128 columns (layer-1 pattern), alone");
    run<32>(out, 1 | 2 | 4 | 8, 4, "  the same + 4 load warps (other columns)");
    run<32>(out, 1 | 2 | 4 | 8, 8, "  the same + 8 load warps (other columns)");
    run<32>(out, 1 | 4 | 16, 4, "ReLU hops only (ld x32 + st x16 + wait + fences), 4 warps");
    run<32>(out, 1 | 2 | 4 | 8 | 16, 4, "N = 32 layer-1 MMAs + 4 warps doing ReLU hops (other columns)");
    run<32>(out, 1 | 2 | 8 | 16, 4, "N = 32 layer-1 MMAs + 4 warps doing ReLU hops IN the accumulator columns");
    run<32>(out, 1 | 2 | 4 | 8 | 16, 8, "N = 32 layer-1 MMAs + 8 warps doing ReLU hops (other columns)");
    run<64>(out, 1 | 2 | 4 | 8, 8, "N = 64, A walking + 8 load warps");
    run<128>(out, 1 | 2 | 4 | 8, 8, "N = 128, A walking + 8 load warps");
    return 0;
}
