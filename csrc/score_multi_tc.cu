// MSE-weight scoring of many models on ONE shared matrix (the pooled dev set), tensor-core path.
//
// This is the scoring half of the fused aggregation (reference: fed_mse_avg, client_trainer.py:115-130,
// which runs a full-batch forward of every selected model on the dev matrix).  The work is
// n_models x rows row-forwards -- quadratic in the number of clients, because the dev set is pooled from
// all of them -- and every model reads the same rows.  score_tc.cu streams the rows once per model; here
// the loop is inverted: a CTA keeps one 128-row tile of X (and of its TF32 residual) resident in shared
// memory and runs every locally owned model over it, so X crosses L2/HBM once per tile instead of once per
// (tile, model).  Models arrive as pre-staged operand images (swizzled UMMA layouts, hi/lo split) that a
// prep kernel builds once per round; the producer thread pulls them with bulk async copies, software
// pipelined against the MMAs and epilogues of the previous model:
//
//   the w1 image is double buffered (image k+2 is requested when epilogue 1 of model k is done), the w4 image
//   of model k is requested as soon as MMA 2 of model k-1 has retired; D1, h2 and D2 are double buffered in
//   TMEM so MMA 1 (k+1) and epilogue 1 (k+1) overlap MMA 2 and epilogue 2 of model k.
//
// Per (tile, model): 48 + 12 tcgen05.mma (3xTF32) or 16 + 4 (TF32), one thread per row for the 27->7->27
// bottleneck and the squared error (x read back from the resident tile with swizzle-aware LDS.128).
#include "tc_common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {
namespace tc {

constexpr int NT_MM = 288;                 // warps 0-3: epilogue 1, warps 4-7: epilogue 2, warp 8: TMA + MMA producer
constexpr int PRODUCER_WARP = 8;
constexpr int MM_TMEM_COLS = 512;
constexpr int C_D1A = 0, C_D1B = 32, C_A2A = 64, C_A2B = 96, C_A2LOA = 128, C_A2LOB = 160, C_D2A = 256, C_D2B = 384;
// small fp32 tables of a model: w2[7][28] | w3[27][8] | b1[32] | b2[8] | b3[32] = 484 floats.  They ride in the
// five unused rows (n = 27..31) of the four K-blocks of the w1 operand image: the tensor core reads them as
// operand rows of output columns nobody looks at, the epilogue reads them as plain floats.  160 floats per block.
constexpr int T_W2 = 0, T_W3 = 196, T_B1 = 412, T_B2 = 444, T_B3 = 452, TBL_FLOATS = 484;
__host__ __device__ constexpr int tbl_off(int t) {          // byte offset of table float t inside a w1 (hi) image
    return (t / 160) * W1BLK + (27 + (t % 160) / 32) * 128 + (t % 32) * 4;
}
static_assert(TBL_FLOATS <= 4 * 160, "tables do not fit in the pad rows");
// operand image of one model in global memory (bytes): w1 hi (+tables) | w1 lo | w4 hi | w4 lo
constexpr int IMG_W1 = 0, IMG_W1LO = 16384, IMG_W4 = 32768, IMG_W4LO = IMG_W4 + 16384;
constexpr int IMG_BYTES = IMG_W4LO + 16384;                // 65,536 B
static_assert(IMG_BYTES == FM_TC_IMAGE_BYTES, "image size drifted from the header");

template <int PREC>
struct MmSmem {
    static constexpr int LO = PREC == 3 ? 1 : 0;
    alignas(1024) unsigned char x[4 * XBLK];
    alignas(1024) unsigned char xlo[LO ? 4 * XBLK : 1024];
    static constexpr int NB1 = 2;                          // w1 image buffers (two hide the image load latency)
    alignas(1024) unsigned char w1[NB1][4 * W1BLK];
    alignas(1024) unsigned char w1lo[NB1][LO ? 4 * W1BLK : 1024];
    alignas(1024) unsigned char w4[128 * 128];
    alignas(1024) unsigned char w4lo[LO ? 128 * 128 : 1024];
    alignas(8) uint64_t bar_x, bar_xfree, bar_w1[2], bar_w4, bar_d1[2], bar_h2[2], bar_d2[2], bar_d2free[2];
    uint32_t tmem_base;
};

// ---- prep: one CTA per slot; owned slots get a dense index and an operand image ----------------
template <int PREC>
__global__ void __launch_bounds__(256) prep_images_kernel(const float* __restrict__ slab, const int* __restrict__ slot_client,
                                                          int first_client, int n_local, int n_slots,
                                                          unsigned char* __restrict__ images, int* __restrict__ owned_slots,
                                                          int* __restrict__ n_owned) {
    __shared__ int s_idx;
    const int slot = blockIdx.x, tid = threadIdx.x;
    const int c = (slot_client ? slot_client[slot] : first_client + slot) - first_client;
    if (c < 0 || c >= n_local) return;
    if (tid == 0) { s_idx = atomicAdd(n_owned, 1); owned_slots[s_idx] = slot; }
    __syncthreads();
    unsigned char* img = images + (size_t)s_idx * IMG_BYTES;
    for (int i = tid; i < IMG_BYTES / 16; i += 256) reinterpret_cast<uint4*>(img)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    const float* row = slab + (size_t)c * P_PAD;
    auto tset = [&](int t, float v) { *reinterpret_cast<float*>(img + IMG_W1 + tbl_off(t)) = v; };
    auto put = [&](int off_hi, int off_lo, int off, float v) {
        if (PREC == 3) {
            const uint32_t hi = trunc_tf32(v);
            *reinterpret_cast<uint32_t*>(img + off_hi + off) = hi;
            *reinterpret_cast<float*>(img + off_lo + off) = v - __uint_as_float(hi);
        } else *reinterpret_cast<uint32_t*>(img + off_hi + off) = cvt_tf32(v);
    };
    for (int p = tid; p < NP; p += 256) {
        const float v = row[p];
        if (p < OFF_B1) { const int n = p / D, k = p - n * D; put(IMG_W1, IMG_W1LO, (k >> 5) * W1BLK + sw128(n, k & 31), v); }
        else if (p < OFF_W2) tset(T_B1 + p - OFF_B1, v);
        else if (p < OFF_B2) { const int q = p - OFF_W2, i = q / H, j = q - i * H; tset(T_W2 + i * 28 + j, v); }
        else if (p < OFF_W3) tset(T_B2 + p - OFF_B2, v);
        else if (p < OFF_B3) { const int q = p - OFF_W3, j = q / Z, i = q - j * Z; tset(T_W3 + j * 8 + i, v); }
        else if (p < OFF_W4) tset(T_B3 + p - OFF_B3, v);
        else if (p < OFF_B4) { const int q = p - OFF_W4, d = q / H, j = q - d * H; put(IMG_W4, IMG_W4LO, sw128(d, j), v); }
        else put(IMG_W4, IMG_W4LO, sw128(p - OFF_B4, 27), v);       // b4 rides in column 27 (h2[27] == 1)
    }
}

template <int PREC>
__global__ void __launch_bounds__(NT_MM, 1)
score_multi_tc_kernel(const __grid_constant__ CUtensorMap xmap, const __grid_constant__ CUtensorMap xlomap,
                      const unsigned char* __restrict__ images, const int* __restrict__ owned_slots,
                      const int* __restrict__ n_owned_ptr, int x_rows, float* __restrict__ out_sse,
                      unsigned long long* __restrict__ dbg) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    using SM = MmSmem<PREC>;
    SM& S = *reinterpret_cast<SM*>(smem_dyn);                // 3xTF32 uses 224 KB of operands: no room for alignment slack
    if ((smem_u32(smem_dyn) & 1023u) != 0u) __trap();         // 128-B swizzle needs a 1024-B aligned base
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_models = *n_owned_ptr;
    const int n_tiles = (x_rows + TROWS - 1) / TROWS;
    if (n_models <= 0 || (int)blockIdx.x >= n_tiles) return;

    if (tid == 0) {
        mbar_init(&S.bar_x, 1); mbar_init(&S.bar_w1[0], 1); mbar_init(&S.bar_w1[1], 1); mbar_init(&S.bar_w4, 1);
        mbar_init(&S.bar_d1[0], 1); mbar_init(&S.bar_d1[1], 1); mbar_init(&S.bar_d2[0], 1); mbar_init(&S.bar_d2[1], 1);
        mbar_init(&S.bar_h2[0], 128); mbar_init(&S.bar_h2[1], 128); mbar_init(&S.bar_xfree, 128);
        mbar_init(&S.bar_d2free[0], 128); mbar_init(&S.bar_d2free[1], 128);
        fence_mbar_init();
    }
    if (warp == PRODUCER_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(MM_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    constexpr uint32_t IDESC1 = idesc_tf32(128, 32), IDESC2 = idesc_tf32(128, 128);
    constexpr uint32_t XBYTES = (PREC == 3 ? 2u : 1u) * 4u * XBLK;
    constexpr uint32_t W1BYTES = (PREC == 3 ? 2u : 1u) * 16384u;
    constexpr uint32_t W4BYTES = (PREC == 3 ? 2u : 1u) * 16384u;
    const int my_tiles = (n_tiles - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;
    const long long total = (long long)my_tiles * n_models;       // (tile, model) pairs of this CTA, in order

    // Software pipeline inside a tile (n_models pairs): the tensor core runs MMA 1 of model k+1 while the
    // epilogue warps finish model k, and the epilogue warps run epilogue 1 of model k+1 while the tensor core
    // runs MMA 2 of model k.  D1, the h2 operand and D2 are all double buffered in TMEM (buffer = pair & 1).
    // The two epilogues run on different warp groups (both groups own the same TMEM lane quarters), one pair apart.
    // Rule kept throughout: no waiter can fall two phases behind the mbarrier it waits on.
    if (warp == PRODUCER_WARP) {
        if (lane == 0) {
            auto load_w1 = [&](long long g) {             // w1 image + tables of pair g
                const unsigned char* img = images + (size_t)(g % n_models) * IMG_BYTES;
                const int wb = (int)(g % SM::NB1);
                mbar_expect_tx(&S.bar_w1[wb], W1BYTES);
                bulk_load(S.w1[wb], img + IMG_W1, 16384, &S.bar_w1[wb]);
                if (PREC == 3) bulk_load(S.w1lo[wb], img + IMG_W1LO, 16384, &S.bar_w1[wb]);
            };
            auto load_w4 = [&](long long g) {
                const unsigned char* img = images + (size_t)(g % n_models) * IMG_BYTES;
                mbar_expect_tx(&S.bar_w4, W4BYTES);
                bulk_load(S.w4, img + IMG_W4, 16384, &S.bar_w4);
                if (PREC == 3) bulk_load(S.w4lo, img + IMG_W4LO, 16384, &S.bar_w4);
            };
            auto load_x = [&](int tile_i) {
                const int row0 = ((int)blockIdx.x + tile_i * (int)gridDim.x) * TROWS;
                mbar_expect_tx(&S.bar_x, XBYTES);
#pragma unroll
                for (int kb = 0; kb < 4; ++kb) {
                    tma_load_2d(S.x + kb * XBLK, &xmap, kb * 32, row0, &S.bar_x);
                    if (PREC == 3) tma_load_2d(S.xlo + kb * XBLK, &xlomap, kb * 32, row0, &S.bar_x);
                }
            };
            const uint64_t dx = smem_desc_sw128(S.x), dxlo = smem_desc_sw128(S.xlo);
            const uint64_t dw1a = smem_desc_sw128(S.w1[0]), dw1b = smem_desc_sw128(S.w1[SM::NB1 - 1]);
            const uint64_t dw1loa = smem_desc_sw128(S.w1lo[0]), dw1lob = smem_desc_sw128(S.w1lo[SM::NB1 - 1]);
            const uint64_t dw4 = smem_desc_sw128(S.w4), dw4lo = smem_desc_sw128(S.w4lo);
            auto issue_mma1 = [&](long long g) {          // needs: x tile resident, w1(g) landed, D1[g&1] free
                const int wb = (int)(g % SM::NB1);
                mbar_wait(&S.bar_w1[wb], (uint32_t)((g / SM::NB1) & 1));
                tc_fence_after();
                const uint32_t d1 = tmem + ((g & 1) ? C_D1B : C_D1A);
                // descriptors differ only in the 14-bit start-address field (units of 16 B): base + constant
                const uint64_t bw1 = wb ? dw1b : dw1a, bw1lo = wb ? dw1lob : dw1loa;
#pragma unroll
                for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint64_t ad = dx + (uint64_t)((kb * XBLK + ks * 32) >> 4);
                        const uint64_t bd = bw1 + (uint64_t)((kb * W1BLK + ks * 32) >> 4);
                        mma_ss(d1, ad, bd, IDESC1, (kb | ks) ? 1u : 0u);
                        if (PREC == 3) {
                            mma_ss(d1, ad, bw1lo + (uint64_t)((kb * W1BLK + ks * 32) >> 4), IDESC1, 1u);
                            mma_ss(d1, dxlo + (uint64_t)((kb * XBLK + ks * 32) >> 4), bd, IDESC1, 1u);
                        }
                    }
                }
                tc_commit(&S.bar_d1[g & 1]);
            };
            long long t_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            long long t_prev = clock64();
            const bool timing = dbg != nullptr && blockIdx.x == 0;
#define MM_TICK(slot) do { if (timing) { const long long now_ = clock64(); t_acc[slot] += now_ - t_prev; t_prev = now_; } } while (0)
            load_x(0); load_w1(0); load_w4(0);
            if (SM::NB1 == 2 && total > 1) load_w1(1);      // second image in flight from the start
            long long g = 0;
            for (int tile_i = 0; tile_i < my_tiles; ++tile_i) {
                mbar_wait(&S.bar_x, (uint32_t)(tile_i & 1));
                issue_mma1(g);
                for (int k = 0; k < n_models; ++k, ++g) {
                    const uint32_t b = (uint32_t)(g & 1), ph2 = (uint32_t)((g >> 1) & 1);
                    MM_TICK(0);
                    mbar_wait(&S.bar_d1[b], ph2);                   // MMA 1 (g) done: its w1 buffer is reusable
                    MM_TICK(1);
                    if (SM::NB1 == 1) { if (g + 1 < total) load_w1(g + 1); }
                    if (g >= 1) {                                   // MMA 2 (g-1) precedes MMA 1 (g) in the pipe: done by now
                        mbar_wait(&S.bar_d2[b ^ 1u], (uint32_t)(((g - 1) >> 1) & 1));
                        load_w4(g);                                 // w4 buffer free again -> image of this pair
                    }
                    if (k + 1 < n_models) issue_mma1(g + 1);        // same x tile; D1[(g+1)&1] was released by epi 1 (g-1)
                    MM_TICK(2);
                    mbar_wait(&S.bar_h2[b], ph2);                   // epilogue 1 (g) done: h2(g) in TMEM, tables[b] free
                    MM_TICK(3);
                    if (SM::NB1 == 2) { if (g + 2 < total) load_w1(g + 2); }   // into the buffer MMA 1 (g) just released
                    mbar_wait(&S.bar_w4, (uint32_t)(g & 1));
                    MM_TICK(4);
                    if (g >= 2) mbar_wait(&S.bar_d2free[b], (uint32_t)(((g - 2) >> 1) & 1));   // epilogue 2 (g-2) has read D2[b]
                    MM_TICK(5);
                    tc_fence_after();
                    const uint32_t d2 = tmem + (b ? C_D2B : C_D2A), a2 = tmem + (b ? C_A2B : C_A2A), a2lo = tmem + (b ? C_A2LOB : C_A2LOA);
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {
                        const uint64_t bd = dw4 + (uint64_t)((ks * 32) >> 4);
                        mma_ts(d2, a2 + ks * 8, bd, IDESC2, ks ? 1u : 0u);
                        if (PREC == 3) {
                            mma_ts(d2, a2 + ks * 8, dw4lo + (uint64_t)((ks * 32) >> 4), IDESC2, 1u);
                            mma_ts(d2, a2lo + ks * 8, bd, IDESC2, 1u);
                        }
                    }
                    tc_commit(&S.bar_d2[b]);                        // completion is awaited one pair later (see above)
                    MM_TICK(6);
                }
                // tile boundary: drain.  Every MMA 1 of this tile has completed; epilogue 2 of the last model still
                // reads x from shared memory, so the refill waits for the epilogue warps' release.
                mbar_wait(&S.bar_xfree, (uint32_t)(tile_i & 1));
                if (tile_i + 1 < my_tiles) load_x(tile_i + 1);
                MM_TICK(7);
            }
            if (timing) for (int i = 0; i < 8; ++i) dbg[i] = (unsigned long long)t_acc[i];
        }
        __syncwarp();
    } else {
        const int r = (warp & 3) * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);
        uint32_t v[32];

        auto epilogue1 = [&](long long g) {        // D1 -> bias/ReLU -> 27->7->27 bottleneck -> h2 (hi/lo) into TMEM
            const uint32_t b = (uint32_t)(g & 1), ph2 = (uint32_t)((g >> 1) & 1);
            // tables sit in the pad rows of the w1 image of this pair (buffer g & 1); they landed with it and are
            // visible through bar_d1: the producer observed bar_w1 before it issued MMA 1 (see note above)
            const unsigned char* TB = S.w1[b];
            auto T1 = [&](int t) -> float { return *reinterpret_cast<const float*>(TB + tbl_off(t)); };
            auto T4 = [&](int t) -> float4 { return *reinterpret_cast<const float4*>(TB + tbl_off(t)); };
            mbar_wait(&S.bar_d1[b], ph2);
            tc_fence_after();
            tmem_ld32(lane_base + (b ? C_D1B : C_D1A), v);
            float h1[28];
#pragma unroll
            for (int j = 0; j < H; ++j) h1[j] = fmaxf(__uint_as_float(v[j]) + T1(T_B1 + j), 0.f);
            h1[27] = 0.f;
            float z[8];
#pragma unroll
            for (int i = 0; i < Z; ++i) {
                float acc = T1(T_B2 + i);
#pragma unroll
                for (int j4 = 0; j4 < 28; j4 += 4) {
                    const float4 w = T4(T_W2 + i * 28 + j4);
                    acc = fmaf(h1[j4], w.x, fmaf(h1[j4 + 1], w.y, fmaf(h1[j4 + 2], w.z, fmaf(h1[j4 + 3], w.w, acc))));
                }
                z[i] = acc;
            }
#pragma unroll
            for (int j = 0; j < H; ++j) {
                const float4 wa = T4(T_W3 + j * 8), wb = T4(T_W3 + j * 8 + 4);
                float acc = T1(T_B3 + j);
                acc = fmaf(z[0], wa.x, fmaf(z[1], wa.y, fmaf(z[2], wa.z, fmaf(z[3], wa.w, acc))));
                acc = fmaf(z[4], wb.x, fmaf(z[5], wb.y, fmaf(z[6], wb.z, acc)));
                const float hv = fmaxf(acc, 0.f);
                if (PREC == 3) { v[j] = trunc_tf32(hv); h1[j] = hv - __uint_as_float(v[j]); }
                else v[j] = cvt_tf32(hv);
            }
            v[27] = __float_as_uint(1.0f);
#pragma unroll
            for (int j = H + 1; j < 32; ++j) v[j] = 0u;
            if (g >= 2) {                                  // MMA 2 (g-2) must have consumed the previous h2 in this buffer
                mbar_wait(&S.bar_d2[b], (uint32_t)(((g - 2) >> 1) & 1));
                tc_fence_after();
            }
            tmem_st32(lane_base + (b ? C_A2B : C_A2A), v);
            if (PREC == 3) {
#pragma unroll
                for (int j = 0; j < H; ++j) v[j] = __float_as_uint(h1[j]);
                v[27] = 0u;
                tmem_st32(lane_base + (b ? C_A2LOB : C_A2LOA), v);
            }
            tc_fence_before();
            mbar_arrive(&S.bar_h2[b]);
        };
        auto epilogue2 = [&](long long g, int k, int row0) {   // D2 (= x_hat incl. b4) minus x from the resident tile
            const uint32_t b = (uint32_t)(g & 1), ph2 = (uint32_t)((g >> 1) & 1);
            mbar_wait(&S.bar_d2[b], ph2);
            tc_fence_after();
            float sse = 0.f;
#pragma unroll
            for (int cb = 0; cb < 4; ++cb) {
                tmem_ld32(lane_base + (b ? C_D2B : C_D2A) + cb * 32, v);
                const unsigned char* xb = S.x + cb * XBLK + r * 128;
#pragma unroll
                for (int q4 = 0; q4 < 8; ++q4) {
                    const float4 xv = *reinterpret_cast<const float4*>(xb + ((q4 ^ (r & 7)) << 4));
                    const int d0 = cb * 32 + q4 * 4;
                    const float e0 = __uint_as_float(v[q4 * 4 + 0]) - xv.x, e1 = __uint_as_float(v[q4 * 4 + 1]) - xv.y;
                    const float e2 = __uint_as_float(v[q4 * 4 + 2]) - xv.z, e3 = __uint_as_float(v[q4 * 4 + 3]) - xv.w;
                    if (d0 + 0 < D) sse = fmaf(e0, e0, sse);
                    if (d0 + 1 < D) sse = fmaf(e1, e1, sse);
                    if (d0 + 2 < D) sse = fmaf(e2, e2, sse);
                    if (d0 + 3 < D) sse = fmaf(e3, e3, sse);
                }
            }
            const int slot = owned_slots[k];
            if (row0 + r < x_rows) out_sse[(size_t)slot * x_rows + row0 + r] = sse;
        };

        long long g = 0;
        if (warp < 4) {                            // group A: epilogue 1 of every pair, in order
            for (; g < total; ++g) epilogue1(g);
        } else {                                   // group B: epilogue 2 of every pair, releases D2 and the x tile
            for (int tile_i = 0; tile_i < my_tiles; ++tile_i) {
                const int row0 = ((int)blockIdx.x + tile_i * (int)gridDim.x) * TROWS;
                mbar_wait(&S.bar_x, (uint32_t)(tile_i & 1));        // acquire the TMA-written tile for generic loads
                for (int k = 0; k < n_models; ++k, ++g) {
                    epilogue2(g, k, row0);
                    tc_fence_before();
                    mbar_arrive(&S.bar_d2free[g & 1]);
                }
                mbar_arrive(&S.bar_xfree);                          // done reading this x tile
            }
        }
    }
    tc_fence_before();
    __syncthreads();
    if (warp == PRODUCER_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(MM_TMEM_COLS));
}

}  // namespace tc
}  // namespace fm

namespace {
typedef CUresult (*EncodeTiledFn2)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int encode_x_map(CUtensorMap* map, const float* base, int64_t rows) {
    static EncodeTiledFn2 fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return (int)cudaErrorNotSupported;
        fn = reinterpret_cast<EncodeTiledFn2>(p);
    }
    const cuuint64_t gdim[2] = {(cuuint64_t)fm::XP, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)fm::XP * sizeof(float)};
    const cuuint32_t box[2] = {32u, (cuuint32_t)fm::tc::TROWS};
    const cuuint32_t estride[2] = {1u, 1u};
    return fn(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estride,
              CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
              CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE) == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}
}  // namespace

extern "C" int fm_score_multi_tc(const FmMultiScoreArgs* a, void* stream) {
    using namespace fm;
    using namespace fm::tc;
    if (a->n_slots <= 0 || a->x_rows <= 0) return 0;
    if (a->precision != 1 && a->precision != 3) return (int)cudaErrorInvalidValue;
    const bool x3 = a->precision == 3;
    if (x3 && !a->X_lo) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t err = cudaMemsetAsync(a->n_owned, 0, sizeof(int), st);
    if (err != cudaSuccess) return (int)err;
    if (x3) prep_images_kernel<3><<<a->n_slots, 256, 0, st>>>(a->slab, a->slot_client, a->first_client, a->n_local, a->n_slots,
                                                               a->images, a->owned_slots, a->n_owned);
    else prep_images_kernel<1><<<a->n_slots, 256, 0, st>>>(a->slab, a->slot_client, a->first_client, a->n_local, a->n_slots,
                                                            a->images, a->owned_slots, a->n_owned);
    CUtensorMap map, lomap;
    int rc = encode_x_map(&map, a->X, a->x_rows);
    if (rc) return rc;
    rc = encode_x_map(&lomap, x3 ? a->X_lo : a->X, a->x_rows);
    if (rc) return rc;
    static int sms = 0;
    if (!sms) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const size_t smem = x3 ? sizeof(MmSmem<3>) : sizeof(MmSmem<1>);
    err = x3 ? cudaFuncSetAttribute(score_multi_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
             : cudaFuncSetAttribute(score_multi_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    const int n_tiles = (int)((a->x_rows + TROWS - 1) / TROWS);
    const int grid = n_tiles < sms ? n_tiles : sms;
    if (x3) score_multi_tc_kernel<3><<<grid, NT_MM, smem, st>>>(map, lomap, a->images, a->owned_slots, a->n_owned, (int)a->x_rows, a->out_sse, a->dbg_cycles);
    else score_multi_tc_kernel<1><<<grid, NT_MM, smem, st>>>(map, lomap, a->images, a->owned_slots, a->n_owned, (int)a->x_rows, a->out_sse, a->dbg_cycles);
    return (int)cudaGetLastError();
}
