// Probe: how fast can ONE SM pull L2-resident data into its shared memory, by mechanism?
//
// The fused scoring kernel (csrc/score_fused_f16.cu) fetches a 36 KB operand image per (tile, model) pair with three
// cp.async.bulk copies.  With its layer-1 protocol isolated (scripts/mma_rate_probe.cu) a 16 KB image per 24 MMAs held the
// MMA thread to ~465 cycles per group whatever the ring depth, i.e. ~35 B / clk, on one CTA as on 148.  This probe times
// the copy alone: 2.3 MB (62 images) re-read round and round from L2 into a ring of shared-memory stages by
//   0  one thread, cp.async.bulk, 16 KB per copy           1  one thread, cp.async.bulk, 4 x 4 KB per stage
//   2  two threads (two warps), 8 KB bulk copies each        3  four threads (four warps), 4 KB bulk copies each
//   4  128 threads, ld.global.v4 (L2 hit) + st.shared.v4      5  128 threads, cp.async 16 B (LDGSTS), groups of 8 per thread
//   6  256 threads, cp.async 16 B
// and prints bytes per clock per SM for a grid of 1 and of 148 CTAs (all CTAs read the same images, as the kernel does).
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Icsrc scripts/l2_to_smem_probe.cu -o build/l2_to_smem_probe
#include <cstdio>
#include <cstdlib>
#include "tc_common.cuh"

using namespace fm;
using namespace fm::tc;

constexpr int STAGE = 16384, NST = 4, REPS = 600, IMG_STRIDE = 36864, N_IMG = 62;

struct CSmem {
    alignas(1024) unsigned char st[NST][STAGE];
    alignas(8) uint64_t full[NST];
};

template <int MODE>
__global__ void __launch_bounds__(256, 1) copy_probe(const unsigned char* __restrict__ g, long long* out, int rot) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    CSmem& S = *reinterpret_cast<CSmem*>(smem_dyn);
    const int tid = threadIdx.x, warp = tid >> 5;
    const int nsrc = MODE == 1 ? 4 : MODE == 2 ? 2 : MODE == 3 ? 4 : 1;      // bulk copies per stage
    if (tid == 0) { for (int i = 0; i < NST; ++i) mbar_init(&S.full[i], nsrc == 4 && MODE == 1 ? 1 : nsrc); fence_mbar_init(); }
    __syncthreads();
    const int first = rot ? (int)blockIdx.x % N_IMG : 0;
    const long long t0 = clock64();
    if (MODE <= 3) {
        // issuing threads: MODE 0 / 1: thread 0; MODE 2: lane 0 of warps 0, 1; MODE 3: lane 0 of warps 0..3
        const int n_thr = MODE == 2 ? 2 : MODE == 3 ? 4 : 1;
        if ((tid & 31) == 0 && warp < n_thr) {
            const int chunk = STAGE / n_thr;
            for (int l = 0; l < REPS; ++l) {
                const int s = l % NST;
                if (l >= NST) mbar_wait(&S.full[s], (uint32_t)((l / NST - 1) & 1));      // previous copy into this stage has landed
                const unsigned char* src = g + (size_t)((first + l) % N_IMG) * IMG_STRIDE + warp * chunk;
                mbar_expect_tx(&S.full[s], (uint32_t)chunk);
                if (MODE == 1) {
                    for (int c = 0; c < 4; ++c) bulk_g2s(S.st[s] + c * 4096, src + c * 4096, 4096u, &S.full[s]);
                } else {
                    bulk_g2s(S.st[s] + warp * chunk, src, (uint32_t)chunk, &S.full[s]);
                }
            }
            for (int l = REPS - NST; l < REPS; ++l) mbar_wait(&S.full[l % NST], (uint32_t)((l / NST) & 1));
        }
    } else if (MODE == 4) {
        if (tid < 128) {
            for (int l = 0; l < REPS; ++l) {
                const uint4* src = reinterpret_cast<const uint4*>(g + (size_t)((first + l) % N_IMG) * IMG_STRIDE);
                uint4* dst = reinterpret_cast<uint4*>(S.st[l % NST]);
                uint4 v[8];
#pragma unroll
                for (int i = 0; i < 8; ++i) v[i] = __ldcg(src + tid + i * 128);
#pragma unroll
                for (int i = 0; i < 8; ++i) dst[tid + i * 128] = v[i];
            }
        }
    } else {
        const int nt = MODE == 5 ? 128 : 256, per = STAGE / 16 / nt;
        if (tid < nt) {
            for (int l = 0; l < REPS; ++l) {
                const unsigned char* src = g + (size_t)((first + l) % N_IMG) * IMG_STRIDE;
                unsigned char* dst = S.st[l % NST];
                for (int i = 0; i < per; ++i) {
                    const int o = (tid + i * nt) * 16;
                    asm volatile("cp.async.cg.shared.global [%0], [%1], 16;" ::"r"(smem_u32(dst + o)), "l"(src + o) : "memory");
                }
                asm volatile("cp.async.commit_group;" ::: "memory");
                asm volatile("cp.async.wait_group %0;" ::"n"(NST - 1) : "memory");
            }
            asm volatile("cp.async.wait_group 0;" ::: "memory");
        }
    }
    __syncthreads();
    if (tid == 0) out[blockIdx.x] = clock64() - t0;
}

template <int MODE>
void run(const unsigned char* g, long long* out, const char* what) {
    cudaFuncSetAttribute(copy_probe<MODE>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)sizeof(CSmem) + 1024);
    for (int rot = 0; rot < 2; ++rot)
        for (int grid : {1, 148}) {
            if (rot && grid == 1) continue;
            copy_probe<MODE><<<grid, 256, sizeof(CSmem) + 1024>>>(g, out, rot);      // warm-up: images into L2
            copy_probe<MODE><<<grid, 256, sizeof(CSmem) + 1024>>>(g, out, rot);
            cudaError_t e = cudaDeviceSynchronize();
            if (e != cudaSuccess) { printf("error: %s\n", cudaGetErrorString(e)); exit(1); }
            long long h[148];
            cudaMemcpy(h, out, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
            long long mx = 0;
            for (int i = 0; i < grid; ++i) mx = h[i] > mx ? h[i] : mx;
            printf("  %-44s grid %3d %s: %6.1f B/clk/SM  (%5.0f cycles per 16 KB)\n", what, grid, rot ? "rotated  " : "lock step",
                   (double)REPS * STAGE / (double)mx, (double)mx / REPS);
        }
}

int main(int argc, char** argv) {
    const int which = argc > 1 ? atoi(argv[1]) : -1;
    unsigned char* g;
    long long* out;
    cudaMalloc(&g, (size_t)N_IMG * IMG_STRIDE);
    cudaMemset(g, 1, (size_t)N_IMG * IMG_STRIDE);
    cudaMalloc(&out, 148 * sizeof(long long));
    if (which < 0 || which == 0) run<0>(g, out, "1 thread, bulk 16 KB");
    if (which < 0 || which == 1) run<1>(g, out, "1 thread, bulk 4 x 4 KB");
    if (which < 0 || which == 2) run<2>(g, out, "2 threads, bulk 8 KB each");
    if (which < 0 || which == 3) run<3>(g, out, "4 threads, bulk 4 KB each");
    if (which < 0 || which == 4) run<4>(g, out, "128 threads, ld.global.v4 + st.shared.v4");
    if (which < 0 || which == 5) run<5>(g, out, "128 threads, cp.async 16 B");
    if (which < 0 || which == 6) run<6>(g, out, "256 threads, cp.async 16 B");
    return 0;
}
