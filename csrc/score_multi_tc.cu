// MSE-weight scoring of many models on ONE shared matrix (the pooled dev set), tensor-core path.
//
// This is the scoring half of the fused aggregation (reference: fed_mse_avg, client_trainer.py:115-130,
// which runs a full-batch forward of every selected model on the dev matrix).  The work is
// n_models x rows row-forwards -- quadratic in the number of clients, because the dev set is pooled from
// all of them -- and every model reads the same rows.  score_tc.cu streams the rows once per model; here
// the loop is inverted: a CTA keeps one 128-row tile of X resident ON CHIP and runs a chunk of the locally
// owned models over it, so X crosses L2/HBM once per unit instead of once per (tile, model).
//
// Where the operands live (v5):
//   * X is the A operand of GEMM 1 for every model of the unit, so it is parked in TENSOR MEMORY: the epilogue
//     warps split the fp32 tile into its TF32 head and the exact fp32 remainder (x = hi + lo) and store both
//     with tcgen05.st (2 x 128 columns).  GEMM 1 then runs in TS mode (A from TMEM) at the tensor-core floor
//     (128 x 32 x 8 in ~16 cycles) instead of re-reading 4 KB of A from shared memory per MMA (~66 cycles
//     measured in v4, where the producer's MMA-1 issue was the bottleneck of the kernel).
//   * The fp32 tile passes through ONE shared-memory staging buffer (TMA, 128-B swizzle): the epilogue-1 warps park
//     it in TMEM, each epilogue-2 thread copies its row into registers for the squared error, and the buffer is
//     immediately refilled with the next unit's tile.
//   * Models arrive as pre-staged operand images (swizzled UMMA layouts, hi/lo split) that a prep kernel builds once
//     per round; a loader warp pulls them with bulk async copies (the MMA-issuing warp only waits and issues: its
//     instruction stream, not the tensor pipe, was the bottleneck when it also loaded).  Both images are double
//     buffered: w1 of model k+2 is requested when epilogue 1 of model k is done (its bias / bottleneck tables ride in
//     the pad rows of that image), w4 of model k+1 when MMA 2 of model k-1 has retired.
//   * D1 is double buffered in TMEM; epilogue 1 writes the TF32 head of h2 back into the D1 columns it just
//     read (the tensor pipe runs MMAs in issue order, and MMA 1 (k+2) is issued after MMA 2 (k)), the
//     remainder of h2 goes to its own double-buffered columns; D2 is single buffered and released by
//     epilogue 2 as soon as it sits in registers.
//
// TMEM map (512 columns): X hi 0..127 | X lo 128..255 | D1 / h2 hi 256+32b | h2 lo 320+32b | D2 384..511.
//
// Work units: (tile, chunk of models).  The host picks the number of chunks per tile so that the units
// divide evenly over the SMs (655 tiles over 148 SMs would leave 11 % of the machine idle in the last wave).
//
// Per (tile, model): 45 + 12 tcgen05.mma (3xTF32) or 15 + 4 (TF32), one thread per row for the 27->7->27
// bottleneck and the squared error (x read back from the resident tile with swizzle-aware LDS.128).
#include <type_traits>
#include "tc_common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {
namespace tc {

constexpr int NT_MM = 320;                 // warps 0-3: epilogue 1 (+ X -> TMEM), 4-7: epilogue 2, 8: MMA issue, 9: TMA / bulk loads
constexpr int PRODUCER_WARP = 8, LOADER_WARP = 9;
constexpr int MM_TMEM_COLS = 512;
constexpr int C_XHI = 0, C_XLO = 128, C_D1 = 256, C_A2LO = 320, C_D2 = 384;
constexpr int MM_MAX_CHUNKS = 4;
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
    alignas(1024) unsigned char x[4 * XBLK];               // staging of the fp32 tile: holds the NEXT unit's rows most of the time
    alignas(1024) unsigned char w1[2][4 * W1BLK];
    alignas(1024) unsigned char w1lo[2][LO ? 4 * W1BLK : 1024];
    alignas(1024) unsigned char w4[2][128 * 128];
    alignas(1024) unsigned char w4lo[2][LO ? 128 * 128 : 1024];
    alignas(8) uint64_t bar_x, bar_xfree, bar_xt, bar_w1[2], bar_w4[2], bar_d1[2], bar_h2[2], bar_d2[2], bar_d2free;
    uint32_t tmem_base;
};

// cursor over the (unit, model) pairs of one CTA, in processing order; every role walks the same sequence
struct PairCursor {
    int i, m, m1, tile;
    bool ok;
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

template <int PREC, bool TIMING>
__global__ void __launch_bounds__(NT_MM, 1)
score_multi_tc_kernel(const __grid_constant__ CUtensorMap xmap, const unsigned char* __restrict__ images,
                      const int* __restrict__ owned_slots, const int* __restrict__ n_owned_ptr, int x_rows,
                      int n_chunks, float* __restrict__ out_sse, unsigned long long* __restrict__ dbg) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    using SM = MmSmem<PREC>;
    SM& S = *reinterpret_cast<SM*>(smem_dyn);                // 3xTF32 uses 224 KB of operands: no room for alignment slack
    if ((smem_u32(smem_dyn) & 1023u) != 0u) __trap();         // 128-B swizzle needs a 1024-B aligned base
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_models = *n_owned_ptr;
    const int n_tiles = (x_rows + TROWS - 1) / TROWS;
    const int n_units = n_tiles * n_chunks;
    if (n_models <= 0 || (int)blockIdx.x >= n_units) return;
    const int my_units = (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x;

    // unit u = (tile u / n_chunks, models [ch * n / n_chunks, (ch + 1) * n / n_chunks)); empty chunks are skipped
    auto next_unit = [&](PairCursor& c) {
        for (;;) {
            if (++c.i >= my_units) { c.ok = false; return; }
            const int u = (int)blockIdx.x + c.i * (int)gridDim.x, ch = u % n_chunks;
            c.tile = u / n_chunks;
            c.m = (int)((long long)ch * n_models / n_chunks);
            c.m1 = (int)((long long)(ch + 1) * n_models / n_chunks);
            if (c.m < c.m1) { c.ok = true; return; }
        }
    };
    auto first_pair = [&]() { PairCursor c; c.i = -1; c.m = c.m1 = 0; c.tile = 0; c.ok = false; next_unit(c); return c; };
    auto next_pair = [&](PairCursor& c) { if (++c.m >= c.m1) next_unit(c); };

    if (tid == 0) {
        mbar_init(&S.bar_x, 1); mbar_init(&S.bar_w1[0], 1); mbar_init(&S.bar_w1[1], 1);
        mbar_init(&S.bar_w4[0], 1); mbar_init(&S.bar_w4[1], 1); mbar_init(&S.bar_d1[0], 1); mbar_init(&S.bar_d1[1], 1);
        mbar_init(&S.bar_d2[0], 1); mbar_init(&S.bar_d2[1], 1); mbar_init(&S.bar_h2[0], 128); mbar_init(&S.bar_h2[1], 128);
        mbar_init(&S.bar_xfree, 256); mbar_init(&S.bar_xt, 128); mbar_init(&S.bar_d2free, 128);
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
    constexpr uint32_t XBYTES = 4u * XBLK;
    constexpr uint32_t W1BYTES = (PREC == 3 ? 2u : 1u) * 16384u;
    constexpr uint32_t W4BYTES = (PREC == 3 ? 2u : 1u) * 16384u;

    // Pipeline (g = running pair index of this CTA, b = g & 1, ut = running unit index):
    //   tensor pipe order  ... MMA1(g+1) MMA2(g) MMA1(g+2) MMA2(g+1) ...
    //   epilogue 1 (g) starts when MMA1(g) retires, epilogue 2 (g) when MMA2(g) retires; the two run on different warps.
    // Rule kept throughout: a parity wait only works if the waiter observes completion c of a barrier
    // before completion c+1 happens.  For every (barrier, waiter) pair below, completion c+1 depends on work that the
    // waiter itself releases after it has observed c -- by data dependency, not by timing.
    const bool has_work = first_pair().ok;      // false only when every unit of this CTA is an empty chunk
    if (!has_work) {
    } else if (warp == PRODUCER_WARP) {
        // (the single issuing thread is chosen with elect.sync: ptxas then knows exactly one lane is active and emits
        // the tcgen05.mma issue sequence back to back; behind `lane == 0` every UTCHMMA is wrapped in an
        // ELECT / BRA.U.ANY loop, ~10 dependent scalar instructions = ~70 cycles per MMA, which made the issue
        // thread -- not the tensor pipe -- the bottleneck of v1..v4)
        if (elect_one()) {
            // This thread is a single warp with nothing to hide latency behind: every instruction costs ~4+ cycles.
            // It therefore does nothing but wait and issue MMAs (the operand loads live on the loader warp below); the
            // loop body is instantiated once per buffer parity (descriptors become base + immediate) and the
            // instrumentation is compiled out unless the caller asked for it.  v6 spent ~700 instructions per pair
            // here, v7 ~650 (loads included), and both were bound by them rather than by the tensor pipe.
            using P0 = std::integral_constant<int, 0>;
            using P1 = std::integral_constant<int, 1>;
            const uint64_t dw1a = smem_desc_sw128(S.w1[0]), dw1b = smem_desc_sw128(S.w1[1]);
            const uint64_t dw1loa = smem_desc_sw128(S.w1lo[0]), dw1lob = smem_desc_sw128(S.w1lo[1]);
            const uint64_t dw4a = smem_desc_sw128(S.w4[0]), dw4b = smem_desc_sw128(S.w4[1]);
            const uint64_t dw4loa = smem_desc_sw128(S.w4lo[0]), dw4lob = smem_desc_sw128(S.w4lo[1]);
            auto issue_mma1 = [&](auto WB, uint32_t phase) {  // needs: X of the unit in TMEM, w1 landed, D1[wb] free
                constexpr int wb = decltype(WB)::value;
                mbar_wait(&S.bar_w1[wb], phase);
                tc_fence_after();
                const uint32_t d1 = tmem + C_D1 + 32 * wb;
                const uint64_t bw1 = wb ? dw1b : dw1a, bw1lo = wb ? dw1lob : dw1loa;
#pragma unroll
                for (int kk = 0; kk < 15; ++kk) {         // K = 115 -> 15 steps of 8 (columns 115..119 are zero)
                    const int kb = kk >> 2, ks = kk & 3;
                    const uint64_t bd = bw1 + (uint64_t)((kb * W1BLK + ks * 32) >> 4);
                    mma_ts(d1, tmem + C_XHI + kk * 8, bd, IDESC1, kk ? 1u : 0u);
                    if (PREC == 3) {
                        mma_ts(d1, tmem + C_XHI + kk * 8, bw1lo + (uint64_t)((kb * W1BLK + ks * 32) >> 4), IDESC1, 1u);
                        mma_ts(d1, tmem + C_XLO + kk * 8, bd, IDESC1, 1u);
                    }
                }
                tc_commit(&S.bar_d1[wb]);
            };
            long long t_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};
            long long t_prev = TIMING ? clock64() : 0;
            const bool timing = TIMING && blockIdx.x == 0;
#define MM_TICK(slot) do { if (TIMING && timing) { const long long now_ = clock64(); t_acc[slot] += now_ - t_prev; t_prev = now_; } } while (0)
            // look-ahead state: unit of pair g (u0); existence / unit of pair g+1
            PairCursor cl = first_pair();
            int u0 = cl.i;
            next_pair(cl);
            bool ok1 = cl.ok;
            int u1 = cl.i;
            int g = 0, ut = 0;
            mbar_wait(&S.bar_xt, 0u);
            tc_fence_after();
            issue_mma1(P0{}, 0u);
            auto body = [&](auto PB) -> bool {              // one pair; returns false after the last one
                constexpr int b = decltype(PB)::value;
                using PN = std::integral_constant<int, b ^ 1>;
                const uint32_t ph2 = (uint32_t)((g >> 1) & 1);
                if (ok1 && u1 != u0) {                          // group A parks the next tile in TMEM as soon as MMA 1 (g) retires
                    ++ut;
                    if (TIMING && timing && ut <= 4) dbg[16 + 4 * (ut - 1)] = (unsigned long long)clock64();
                    mbar_wait(&S.bar_xt, (uint32_t)(ut & 1));
                    tc_fence_after();
                    if (TIMING && timing && ut <= 4) dbg[17 + 4 * (ut - 1)] = (unsigned long long)clock64();
                }
                MM_TICK(6);
                // D1[b^1]: epilogue 1 (g-1) is done and MMA 2 (g-1), the reader of the h2 head kept there, was issued before
                if (ok1) issue_mma1(PN{}, (uint32_t)(((g + 1) >> 1) & 1));
                MM_TICK(0);
                mbar_wait(&S.bar_h2[b], ph2);                   // epilogue 1 (g) done: h2(g) in TMEM
                MM_TICK(2);
                mbar_wait(&S.bar_w4[b], ph2);
                MM_TICK(3);
                if (g >= 1) mbar_wait(&S.bar_d2free, (uint32_t)(b ^ 1));   // epilogue 2 (g-1) holds D2 in registers
                MM_TICK(4);
                tc_fence_after();
                const uint32_t d2 = tmem + C_D2, a2 = tmem + C_D1 + 32 * b, a2lo = tmem + C_A2LO + 32 * b;
                const uint64_t bw4 = b ? dw4b : dw4a, bw4lo = b ? dw4lob : dw4loa;
#pragma unroll
                for (int ks = 0; ks < 4; ++ks) {
                    const uint64_t bd = bw4 + (uint64_t)((ks * 32) >> 4);
                    mma_ts(d2, a2 + ks * 8, bd, IDESC2, ks ? 1u : 0u);
                    if (PREC == 3) {
                        mma_ts(d2, a2 + ks * 8, bw4lo + (uint64_t)((ks * 32) >> 4), IDESC2, 1u);
                        mma_ts(d2, a2lo + ks * 8, bd, IDESC2, 1u);
                    }
                }
                tc_commit(&S.bar_d2[b]);       // one barrier per parity: see the loader's wait
                MM_TICK(5);
                const bool more = ok1;
                u0 = u1;
                if (ok1) { next_pair(cl); ok1 = cl.ok; u1 = cl.i; }
                ++g;
                return more;
            };
            for (;;) {
                if (!body(P0{})) break;
                if (!body(P1{})) break;
            }
            if (TIMING && timing) for (int i = 0; i < 8; ++i) dbg[i] = (unsigned long long)t_acc[i];
        }
        __syncwarp();
    } else if (warp == LOADER_WARP) {
        // Operand traffic: the x tile of the next unit and the w1 / w4 images two pairs / one pair ahead.  Every buffer
        // is refilled as soon as the barrier that proves its last reader done has completed; the MMA thread only ever
        // waits for "landed" barriers.
        if (elect_one()) {
            auto load_w1 = [&](int m, int wb) {               // w1 image + tables of a pair
                const unsigned char* img = images + (size_t)m * IMG_BYTES;
                mbar_expect_tx(&S.bar_w1[wb], W1BYTES);
                bulk_load(S.w1[wb], img + IMG_W1, 16384, &S.bar_w1[wb]);
                if (PREC == 3) bulk_load(S.w1lo[wb], img + IMG_W1LO, 16384, &S.bar_w1[wb]);
            };
            auto load_w4 = [&](int m, int wb) {
                const unsigned char* img = images + (size_t)m * IMG_BYTES;
                mbar_expect_tx(&S.bar_w4[wb], W4BYTES);
                bulk_load(S.w4[wb], img + IMG_W4, 16384, &S.bar_w4[wb]);
                if (PREC == 3) bulk_load(S.w4lo[wb], img + IMG_W4LO, 16384, &S.bar_w4[wb]);
            };
            auto load_x = [&](int tile) {
                mbar_expect_tx(&S.bar_x, XBYTES);
#pragma unroll
                for (int kb = 0; kb < 4; ++kb) tma_load_2d(S.x + kb * XBLK, &xmap, kb * 32, tile * TROWS, &S.bar_x);
            };
            // look-ahead state: unit of pair g (u0); existence / unit / model of pairs g+1 and g+2
            PairCursor cl = first_pair();
            PairCursor cx = cl;
            load_x(cx.tile); next_unit(cx);
            int u0 = cl.i;
            load_w1(cl.m, 0); load_w4(cl.m, 0);
            next_pair(cl);
            bool ok1 = cl.ok;
            int u1 = cl.i, m1 = cl.m;
            if (ok1) { load_w1(m1, 1); load_w4(m1, 1); next_pair(cl); }
            bool ok2 = ok1 && cl.ok;
            int u2 = cl.i, m2 = cl.m;
            // the staging buffer is refilled with the next unit's rows once group A has parked the current tile in
            // TMEM and group B holds its rows in registers (bar_xfree); checked when that is known to have happened
            int ut = 0, x_release_pending = 0, x_release_at = 2;
            auto flush_x_release = [&]() {
                mbar_wait(&S.bar_xfree, (uint32_t)(x_release_pending & 1));
                if (cx.ok) { load_x(cx.tile); next_unit(cx); }
                x_release_pending = -1;
            };
            for (int g = 0;; ++g) {
                const int b = g & 1;
                if (g >= 1) {                                   // MMA 2 (g-1) retired -> its w4 buffer takes the image of pair g+1
                    // bar_d2 alternates with the pair parity: the next completion of THIS barrier is MMA 2 (g+1), which
                    // needs the image requested right below -- so this thread cannot miss a phase, whatever the timing
                    mbar_wait(&S.bar_d2[b ^ 1], (uint32_t)(((g - 1) >> 1) & 1));
                    if (ok1) load_w4(m1, b ^ 1);
                    if (x_release_pending >= 0 && g >= x_release_at) flush_x_release();   // epilogue 2 (g-2) ran: B is in this unit
                }
                if (ok1 && u1 != u0) {                          // unit boundary ahead: group A needs the next tile staged
                    if (x_release_pending >= 0) flush_x_release();      // (units shorter than three pairs: not requested yet)
                    ++ut;
                    x_release_pending = ut;
                    x_release_at = g + 3;
                }
                mbar_wait(&S.bar_h2[b], (uint32_t)((g >> 1) & 1));   // epilogue 1 (g) done: w1 buffer b (tables) is free
                if (ok2) load_w1(m2, b);
                if (!ok1) break;
                u0 = u1; ok1 = ok2; u1 = u2; m1 = m2;
                if (ok2) { next_pair(cl); ok2 = cl.ok; u2 = cl.i; m2 = cl.m; }
            }
        }
        __syncwarp();
    } else {
        const int r = (warp & 3) * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)((warp & 3) * 32) << 16);

        // 9 warps put three on one SM sub-partition (16 K registers): 168 per thread at launch.  Group B keeps a
        // 116-float row in registers, group A needs far fewer: the two warpgroups trade registers.
        if (warp < 4) {                            // group A: X -> TMEM per unit, epilogue 1 of every pair
            asm volatile("setmaxnreg.dec.sync.aligned.u32 128;" ::);
            uint32_t v[32];
            float h1[32];
            const bool timing = TIMING && blockIdx.x == 0 && tid == 0;
            long long ta[4] = {0, 0, 0, 0};               // bar_x wait | X -> TMEM | bar_d1 wait | epilogue 1 body
            auto stage_x = [&](int ut) {            // fp32 tile -> TF32 head + exact remainder, one row per thread
                long long t0 = timing ? clock64() : 0;
                mbar_wait(&S.bar_x, (uint32_t)(ut & 1));
                if (timing) { const long long t1 = clock64(); ta[0] += t1 - t0; t0 = t1; }
#pragma unroll
                for (int cb = 0; cb < 4; ++cb) {
                    const unsigned char* xb = S.x + cb * XBLK + r * 128;
#pragma unroll
                    for (int q4 = 0; q4 < 8; ++q4) {
                        const float4 xv = *reinterpret_cast<const float4*>(xb + ((q4 ^ (r & 7)) << 4));
                        const float xs[4] = {xv.x, xv.y, xv.z, xv.w};
#pragma unroll
                        for (int e = 0; e < 4; ++e) {
                            if (PREC == 3) { v[q4 * 4 + e] = trunc_tf32(xs[e]); h1[q4 * 4 + e] = xs[e] - __uint_as_float(v[q4 * 4 + e]); }
                            else v[q4 * 4 + e] = cvt_tf32(xs[e]);
                        }
                    }
                    tmem_st32(lane_base + C_XHI + cb * 32, v);
                    if (PREC == 3) {
#pragma unroll
                        for (int j = 0; j < 32; ++j) v[j] = __float_as_uint(h1[j]);
                        tmem_st32(lane_base + C_XLO + cb * 32, v);
                    }
                }
                tc_fence_before();
                mbar_arrive(&S.bar_xt);
                mbar_arrive(&S.bar_xfree);              // this half of the CTA is done with the staged tile
                if (timing) ta[1] += clock64() - t0;
            };
            auto epilogue1 = [&](int g, int stage_ut) {   // D1 -> bias/ReLU -> 27->7->27 bottleneck -> h2 (hi/lo) into TMEM
                const uint32_t b = (uint32_t)(g & 1), ph2 = (uint32_t)((g >> 1) & 1);
                // tables sit in the pad rows of the w1 image of this pair (buffer b); they landed with it and are
                // visible through bar_d1: the producer observed bar_w1 before it issued MMA 1
                const unsigned char* TB = S.w1[b];
                auto T1 = [&](int t) -> float { return *reinterpret_cast<const float*>(TB + tbl_off(t)); };
                auto T4 = [&](int t) -> float4 { return *reinterpret_cast<const float4*>(TB + tbl_off(t)); };
                long long t0 = timing ? clock64() : 0;
                mbar_wait(&S.bar_d1[b], ph2);           // MMA 1 (g) retired -- and with it every MMA issued before it
                tc_fence_after();
                if (timing) { const long long t1 = clock64(); ta[2] += t1 - t0; t0 = t1; }
                // last pair of a unit: MMA 1 (g) was the last reader of the old X, so the next tile goes to TMEM first
                // (the producer issues MMA 1 (g+1) behind it while this epilogue runs)
                if (stage_ut >= 0) {
                    if (timing && stage_ut <= 4) dbg[18 + 4 * (stage_ut - 1)] = (unsigned long long)clock64();
                    stage_x(stage_ut);
                    if (timing) t0 = clock64();
                    if (timing && stage_ut <= 4) dbg[19 + 4 * (stage_ut - 1)] = (unsigned long long)t0;
                }
                tmem_ld32(lane_base + C_D1 + 32 * b, v);
#pragma unroll
                for (int j = 0; j < H; ++j) h1[j] = fmaxf(__uint_as_float(v[j]) + T1(T_B1 + j), 0.f);
                h1[27] = 0.f;
                // 27 -> 7 -> 27 with packed FMAs (FFMA2): even / odd partial sums, 247 issue slots instead of 385
                float z[8];
#pragma unroll
                for (int i = 0; i < Z; ++i) {
                    float2 acc = make_float2(T1(T_B2 + i), 0.f);
#pragma unroll
                    for (int j4 = 0; j4 < 28; j4 += 4) {
                        const float4 w = T4(T_W2 + i * 28 + j4);
                        acc = ffma2(make_float2(h1[j4], h1[j4 + 1]), make_float2(w.x, w.y), acc);
                        acc = ffma2(make_float2(h1[j4 + 2], h1[j4 + 3]), make_float2(w.z, w.w), acc);
                    }
                    z[i] = acc.x + acc.y;
                }
                z[7] = 0.f;
#pragma unroll
                for (int j = 0; j < H; ++j) {
                    const float4 wa = T4(T_W3 + j * 8), wb = T4(T_W3 + j * 8 + 4);     // w3[j][7] is a zero pad
                    float2 acc = make_float2(T1(T_B3 + j), 0.f);
                    acc = ffma2(make_float2(z[0], z[1]), make_float2(wa.x, wa.y), acc);
                    acc = ffma2(make_float2(z[2], z[3]), make_float2(wa.z, wa.w), acc);
                    acc = ffma2(make_float2(z[4], z[5]), make_float2(wb.x, wb.y), acc);
                    acc = ffma2(make_float2(z[6], z[7]), make_float2(wb.z, wb.w), acc);
                    const float hv = fmaxf(acc.x + acc.y, 0.f);
                    if (PREC == 3) { v[j] = trunc_tf32(hv); h1[j] = hv - __uint_as_float(v[j]); }
                    else v[j] = cvt_tf32(hv);
                }
                v[27] = __float_as_uint(1.0f);
#pragma unroll
                for (int j = H + 1; j < 32; ++j) v[j] = 0u;
                // h2 hi replaces D1[b] in place; h2 lo has its own columns.  MMA 2 (g-2), the last reader of both,
                // was issued before MMA 1 (g) and therefore retired before bar_d1 completed.
                tmem_st32(lane_base + C_D1 + 32 * b, v);
                if (PREC == 3) {
#pragma unroll
                    for (int j = 0; j < H; ++j) v[j] = __float_as_uint(h1[j]);
                    v[27] = 0u;
                    tmem_st32(lane_base + C_A2LO + 32 * b, v);
                }
                tc_fence_before();
                mbar_arrive(&S.bar_h2[b]);
                if (timing) ta[3] += clock64() - t0;
            };
            PairCursor c = first_pair();
            int g = 0, ut = 0;
            stage_x(0);
            while (c.ok) {
                PairCursor cn = c;
                next_pair(cn);
                epilogue1(g, (cn.ok && cn.i != c.i) ? ++ut : -1);
                c = cn;
                ++g;
            }
            if (timing) for (int i = 0; i < 4; ++i) dbg[8 + i] = (unsigned long long)ta[i];
        } else {                                   // group B: epilogue 2 of every pair, releases D2 and the x tile
            asm volatile("setmaxnreg.inc.sync.aligned.u32 208;" ::);
            uint32_t v0[32], v1[32];
            float xr[116];                          // this thread's row of the tile, kept in registers for the whole unit
            PairCursor c = first_pair();
            int g = 0, ut = 0;
            bool fresh = true;
            const bool timing = TIMING && blockIdx.x == 0 && tid == 128;
            long long tb[4] = {0, 0, 0, 0};               // unit entry (x -> registers) | bar_d2 wait | until D2 released | rest
            while (c.ok) {
                long long t0 = timing ? clock64() : 0;
                if (fresh) {
                    mbar_wait(&S.bar_x, (uint32_t)(ut & 1));            // acquire the TMA-written tile
#pragma unroll
                    for (int q = 0; q < 29; ++q) {
                        const float4 xv = *reinterpret_cast<const float4*>(S.x + (q >> 3) * XBLK + r * 128 + (((q & 7) ^ (r & 7)) << 4));
                        xr[q * 4 + 0] = xv.x; xr[q * 4 + 1] = xv.y; xr[q * 4 + 2] = xv.z; xr[q * 4 + 3] = xv.w;
                    }
                    mbar_arrive(&S.bar_xfree);
                    fresh = false;
                }
                if (timing) { const long long t1 = clock64(); tb[0] += t1 - t0; t0 = t1; }
                mbar_wait(&S.bar_d2[g & 1], (uint32_t)((g >> 1) & 1));
                tc_fence_after();
                if (timing) { const long long t1 = clock64(); tb[1] += t1 - t0; t0 = t1; }
                float sse = 0.f;
                auto acc_block = [&](const uint32_t (&v)[32], int cb) {
#pragma unroll
                    for (int j = 0; j < 32; ++j) {
                        const int d = cb * 32 + j;
                        if (d < D) { const float e = __uint_as_float(v[j]) - xr[d]; sse = fmaf(e, e, sse); }
                    }
                };
                tmem_ld64(lane_base + C_D2, v0, v1);
                acc_block(v0, 0); acc_block(v1, 1);
                tmem_ld64(lane_base + C_D2 + 64, v0, v1);
                tc_fence_before();
                mbar_arrive(&S.bar_d2free);                   // D2 (= x_hat incl. b4) is in registers: MMA 2 (g+1) may overwrite it
                if (timing) { const long long t1 = clock64(); tb[2] += t1 - t0; t0 = t1; }
                acc_block(v0, 2); acc_block(v1, 3);
                const int row = c.tile * TROWS + r;
                if (row < x_rows) out_sse[(size_t)owned_slots[c.m] * x_rows + row] = sse;
                PairCursor cn = c;
                next_pair(cn);
                if (!cn.ok || cn.i != c.i) { ++ut; fresh = true; }
                c = cn;
                ++g;
                if (timing) tb[3] += clock64() - t0;
            }
            if (timing) for (int i = 0; i < 4; ++i) dbg[12 + i] = (unsigned long long)tb[i];
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
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t err = cudaMemsetAsync(a->n_owned, 0, sizeof(int), st);
    if (err != cudaSuccess) return (int)err;
    if (x3) prep_images_kernel<3><<<a->n_slots, 256, 0, st>>>(a->slab, a->slot_client, a->first_client, a->n_local, a->n_slots,
                                                               a->images, a->owned_slots, a->n_owned);
    else prep_images_kernel<1><<<a->n_slots, 256, 0, st>>>(a->slab, a->slot_client, a->first_client, a->n_local, a->n_slots,
                                                            a->images, a->owned_slots, a->n_owned);
    CUtensorMap map;
    int rc = encode_x_map(&map, a->X, a->x_rows);
    if (rc) return rc;
    static int sms = 0;
    if (!sms) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const size_t smem = x3 ? sizeof(MmSmem<3>) : sizeof(MmSmem<1>);
    using KernelFn = void (*)(const CUtensorMap, const unsigned char*, const int*, const int*, int, int, float*, unsigned long long*);
    const bool timed = a->dbg_cycles != nullptr;
    KernelFn kern = x3 ? (timed ? (KernelFn)score_multi_tc_kernel<3, true> : (KernelFn)score_multi_tc_kernel<3, false>)
                       : (timed ? (KernelFn)score_multi_tc_kernel<1, true> : (KernelFn)score_multi_tc_kernel<1, false>);
    err = cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    // chunks of models per tile: balance the units over the SMs.  Cost model: waves x (models per unit + ~3 pairs
    // of fill/drain per unit); the number of owned models is only known on the device, n_local bounds it.
    const int n_tiles = (int)((a->x_rows + TROWS - 1) / TROWS);
    const int m_est = a->n_slots < a->n_local ? a->n_slots : (a->n_local > 0 ? a->n_local : 1);
    int n_chunks = 1;
    double best = 1e30;
    for (int c = 1; c <= MM_MAX_CHUNKS && c <= (m_est > 0 ? m_est : 1); ++c) {
        const long long units = (long long)n_tiles * c;
        const long long g = units < sms ? units : sms;
        const double cost = (double)((units + g - 1) / g) * ((double)m_est / c + 3.0);
        if (cost < best * 0.98) { best = cost; n_chunks = c; }
    }
    const long long units = (long long)n_tiles * n_chunks;
    const int grid = (int)(units < sms ? units : sms);
    kern<<<grid, NT_MM, smem, st>>>(map, a->images, a->owned_slots, a->n_owned, (int)a->x_rows, n_chunks, a->out_sse, a->dbg_cycles);
    return (int)cudaGetLastError();
}
