// MSE-weight scoring of many models on ONE shared matrix (the pooled dev set), second generation:
// every layer on the tensor core, fp16 hi/lo operands, per-model MSE finished inside the kernel.
//
// Reference: fed_mse_avg, src/Trainer/client_trainer.py:115-130 -- a full-batch forward of every selected model on
// the dev matrix, s_j = 1 / MSE_j.  The work is n_models x rows row-forwards, quadratic in the client count (the dev
// set is pooled from all clients), so it is the one phase of a round whose per-GPU cost grows with the number of
// GPUs; score_multi_tc.cu (3xTF32) spent 1.6 k tensor cycles and 1.6 k CUDA-core cycles per (128-row tile, model).
//
// What changed against score_multi_tc.cu:
//  * fp16 operands instead of 3xTF32, fp32 accumulation in TMEM (kind::f16: twice the TF32 rate, K = 16 per
//    instruction).  Weights and the input tile are split as v = hi + lo with hi = fp16(v), lo = fp16(v - hi) (22 mantissa
//    bits together, |v| saturated at the fp16 range first); GEMM 1 is the three products hi*hi + hi*lo + lo*hi.  The
//    hidden activations h1 / h2 travel as ONE fp16: their rounding error (2^-11, unbiased, different in every row)
//    averages out over the 10^5..10^6 rows of an MSE, whereas a rounding of the weights would be the same in every row.
//    GEMMs 1b / 2 are therefore two products (h * W_hi + h * W_lo): 24 + 4 + 4 MMAs (~0.74 k tensor cycles) per pair
//    instead of 45 + 12 (~1.5 k), and the epilogues between the GEMMs are one convert per pair of values.
//  * The 27 -> 7 -> 27 bottleneck left the CUDA cores.  There is no activation between the two inner layers, so
//    h2 = relu(W3 (W2 h1 + b2) + b3) = relu(W32 h1 + b32) with W32 = W3 W2 (27 x 27, rank 7) and b32 = W3 b2 + b3,
//    formed once per model by the prep kernel: ONE extra small GEMM (M128 N32 K32) between the two big ones.
//  * Biases ride in the GEMMs: the parked tile gets a constant-one column (k = 115), W1 carries b1 in that column
//    and a pass-through row that makes h1[27] = 1, W32 carries b32 in column 27 and passes the one on, W4 carries
//    b4 in column 27.  The epilogues between the GEMMs are ReLU + fp16 split only, no tables, no FMAs.
//  * D2 (the reconstruction, 128 columns) is double buffered in TMEM and read by EIGHT epilogue warps: with one buffer
//    the loop MMA2 -> squared error -> release -> next MMA2 set the pace of the kernel (ncu of the first two versions:
//    1.7 k cycles per pair, tensor pipe 55 % active).  D1 / D1b are double buffered, the parked tile is single buffered
//    in exchange (a unit boundary costs one parking, ~3 % at 62 models per unit).
//  * No [models x rows] SSE matrix: every squared-error thread adds its row's error to its own running sum per model
//    (a fire-and-forget RED to a slot nobody else writes); after a grid-wide barrier the CTAs add those sums in a fixed
//    order to the per-model MSE (deterministic, replaces chunk_mean).
//
// TMEM map (512 columns): X hi 0..63, lo 64..127 | D1/h1[r] 128+32r | D1b/h2[r] 192+32r (r = pair mod 2) | D2[b] 256+128b.
// Warp roles: 0-3 epilogues 1a/1b + tile parking, 4-11 epilogue 2 (squared error; two warps per TMEM lane quarter, half
// of the output columns each), 12 / 13 / 14 MMA issue (one thread per GEMM stage), 15 weight images + X tiles (TMA).
//
// Tensor-pipe issue order in steady state:  ... MMA1(g+2)  MMA1b(g+1)  MMA2(g)  MMA1(g+3) ...
// TMA multicast is not used; the operand images and the dev matrix live in this GPU's memory.
#include <cuda_fp16.h>
#include <type_traits>
#include "tc_common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {
namespace tcf {

using namespace fm::tc;

constexpr int NT_F = 512;                                 // 16 warps, 4 per SM sub-partition: 128 registers per thread
constexpr int MMA1_WARP = 12, MMA1B_WARP = 13, MMA2_WARP = 14, LOAD_WARP = 15;
constexpr int F_TMEM_COLS = 512;
constexpr int NB = 2;                                     // ring depth of the D1 / D1b accumulators (pairs in flight per stage)
constexpr int C_X = 0, C_XLO = 64, C_D1 = 128, C_D1B = C_D1 + 32 * NB, C_D2 = 256;   // D1[r] = 128 + 32 r, D1b[r] = 192 + 32 r, D2[b] = 256 + 128 b
// operand image of one model (bytes), K-major.  W1 (K = 128): 128-B swizzle, two K-blocks of 64 fp16 per row.
// W32 and W4 (K = 32): 64-B swizzle, one 64-byte row per output feature -- no padding travels.  The images are what
// every SM pulls from L2 for every (tile, model) pair, and at ~42 B / clk / SM of L2 bandwidth the 56 KB of the first
// layout (128-B rows everywhere, K padded to 64) cost 1.35 k cycles per pair, more than the 0.93 k of tensor work.
// ONE commit per group of MMAs: "layer 1 / bottleneck / output GEMM of pair q has retired" completes bar_m1 / bar_m1b /
// bar_m2[q % RB], and every consumer of that event waits on it with parity (q / RB) & 1 -- the epilogue warps (accumulator
// ready), the loader (image stage free) and the upstream MMA thread (accumulator slot free).  A commit costs its thread
// ~50-100 cycles, and the bottleneck / output threads used to issue three each for four MMAs.  RB is a multiple of both
// ring depths; no consumer can fall RB pairs behind its producer (each of them gates a ring of depth <= 3).
constexpr int RB = 6;
constexpr int NW = 3;                       // shared-memory stages per image kind
constexpr int W1_BYTES = 2 * 4096;          // [32 n][128 k]  = two K-blocks of 32 rows x 128 B
constexpr int W32_BYTES = 32 * 64;          // [32 n][32 k]
constexpr int W4_BYTES = 128 * 64;          // [128 n][32 k]
constexpr int IMG_W1 = 0, IMG_W1LO = IMG_W1 + W1_BYTES, IMG_W32 = IMG_W1LO + W1_BYTES, IMG_W32LO = IMG_W32 + W32_BYTES,
              IMG_W4 = IMG_W32LO + W32_BYTES, IMG_W4LO = IMG_W4 + W4_BYTES, IMG_F_BYTES = IMG_W4LO + W4_BYTES;   // 36,864
static_assert(IMG_F_BYTES <= FM_TC_IMAGE_BYTES, "the f16 image must fit the image slot of the header");
constexpr float F16_MAX = 65504.f;
// debug cycle slots (CTA 0): issuing threads 1 / 1b / 2: waits (tile or epilogue, image, free buffer) and issue;
// parking+epilogue warp 0: park, e1, e1b, idle polling; squared-error warp 4: wait D2, until D2 released, rest, unit entry
enum { TK_1_XT = 0, TK_1_W, TK_1_FREE, TK_1_ISSUE, TK_1B_H, TK_1B_W, TK_1B_FREE, TK_1B_ISSUE, TK_2_H, TK_2_W, TK_2_FREE, TK_2_ISSUE,
       TK_A_PARK, TK_A_E1, TK_A_E1B, TK_A_IDLE, TK_B_WAIT, TK_B_HOLD, TK_B_REST, TK_B_ENTRY, TK_TOTAL, TK_N };
constexpr int TRACE_N = 96, TRACE_BASE = 128;      // TIMING: time line of the first TRACE_N pairs of CTA 0, 13 events each
#define FT(slot) do { if (TIMING && tk_on) { const long long now_ = clock64(); tk[slot] += now_ - tk_t; tk_t = now_; } } while (0)

struct FSmem {
    alignas(1024) unsigned char x[4 * XBLK];                 // fp32 staging of one 128-row tile (TMA, 128-B swizzle)
    alignas(1024) unsigned char w1[NW][2 * W1_BYTES];        // hi | lo
    alignas(1024) unsigned char w32[NW][2 * W32_BYTES];
    alignas(1024) unsigned char w4[NW][2 * W4_BYTES];
    alignas(8) uint64_t bar_x, bar_xfree, bar_xt, bar_w1[NW], bar_w32[NW], bar_w4[NW], bar_m1[RB], bar_m1b[RB], bar_m2[RB], bar_h1[NB],
        bar_h2[NB], bar_d2free[2];
    uint32_t tmem_base;
    double red[16];
};

// byte offset of fp16 element (row, k) inside a [rows x 64 fp16] K-block with the 128-B swizzle
__host__ __device__ __forceinline__ int sw128h(int row, int k) {
    return row * 128 + ((((k >> 3) ^ (row & 7)) << 4) | ((k & 7) << 1));
}
// the same for a [rows x 32 fp16] operand with the 64-B swizzle: 16-byte chunk c of row r sits at chunk c ^ ((r >> 1) & 3)
__host__ __device__ __forceinline__ int sw64h(int row, int k) {
    return row * 64 + ((((k >> 3) ^ ((row >> 1) & 3)) << 4) | ((k & 7) << 1));
}
// K-major, 64-byte swizzle shared-memory matrix descriptor: 8-row atoms of 512 B (SBO), layout type SWIZZLE_64B = 4
__device__ __forceinline__ uint64_t smem_desc_sw64(const void* p) {
    const uint32_t a = smem_u32(p);
    uint64_t d = 0;
    d |= (uint64_t)((a & 0x3FFFFu) >> 4);
    d |= (uint64_t)1 << 16;
    d |= (uint64_t)(512 >> 4) << 32;
    d |= (uint64_t)1 << 46;
    d |= (uint64_t)4 << 61;
    return d;
}
__host__ __device__ constexpr uint32_t idesc_f16(int M, int N) {       // D = F32, A = B = F16, both K-major, dense
    return (1u << 4) | ((uint32_t)(N >> 3) << 17) | ((uint32_t)(M >> 4) << 24);
}
__device__ __forceinline__ void mma_ts_f16(uint32_t d_tmem, uint32_t a_tmem, uint64_t bdesc, uint32_t idesc, uint32_t accum) {
    asm volatile("{\n\t.reg .pred p;\n\tsetp.ne.b32 p, %4, 0;\n\t"
                 "tcgen05.mma.cta_group::1.kind::f16 [%0], [%1], %2, %3, {%5, %5, %5, %5}, p;\n\t}\n"
                 ::"r"(d_tmem), "r"(a_tmem), "l"(bdesc), "r"(idesc), "r"(accum), "r"(0u) : "memory");
}
__device__ __forceinline__ void tmem_st16(uint32_t taddr, const uint32_t (&v)[16]) {
    asm volatile("tcgen05.st.sync.aligned.32x32b.x16.b32 [%0], "
                 "{%1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15, %16};\n"
                 ::"r"(taddr), "r"(v[0]), "r"(v[1]), "r"(v[2]), "r"(v[3]), "r"(v[4]), "r"(v[5]), "r"(v[6]), "r"(v[7]),
                   "r"(v[8]), "r"(v[9]), "r"(v[10]), "r"(v[11]), "r"(v[12]), "r"(v[13]), "r"(v[14]), "r"(v[15])
                 : "memory");
}
__device__ __forceinline__ void tmem_st_wait() { asm volatile("tcgen05.wait::st.sync.aligned;" ::: "memory"); }
__device__ __forceinline__ bool mbar_test(uint64_t* bar, uint32_t parity) {
    uint32_t ok;
    asm volatile("{\n\t.reg .pred p;\n\tmbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n\tselp.b32 %0, 1, 0, p;\n\t}\n"
                 : "=r"(ok) : "r"(smem_u32(bar)), "r"(parity) : "memory");
    return ok != 0;
}
// (a, b) -> packed fp16 heads (a in the low half: element k of an A row sits in bits [15:0] of column k / 2 when k is
// even -- verified on the device: the other order is off by 1e-3) and packed fp16 remainders.  The converts saturate at +-65504, so the remainder absorbs what the head cannot hold.
__device__ __forceinline__ uint32_t pack_f16x2_sat(float lo_half, float hi_half) {
    uint32_t r;
    asm("cvt.rn.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_half), "f"(lo_half));
    return r;
}
__device__ __forceinline__ uint32_t pack_f16x2_relu_sat(float lo_half, float hi_half) {
    uint32_t r;
    asm("cvt.rn.relu.satfinite.f16x2.f32 %0, %1, %2;" : "=r"(r) : "f"(hi_half), "f"(lo_half));
    return r;
}
__device__ __forceinline__ void split2(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_f16x2_sat(a, b);
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&hi));      // f.x = low half
    lo = pack_f16x2_sat(a - f.x, b - f.y);
}
// the same for relu(a), relu(b): one convert does ReLU + saturation + packing
__device__ __forceinline__ void split2_relu(float a, float b, uint32_t& hi, uint32_t& lo) {
    hi = pack_f16x2_relu_sat(a, b);
    const float2 f = __half22float2(*reinterpret_cast<const __half2*>(&hi));
    const float ra = fmaxf(a, 0.f), rb = fmaxf(b, 0.f);
    lo = pack_f16x2_sat(ra - f.x, rb - f.y);
}
__device__ __forceinline__ float2 fadd2(float2 a, float2 b) {
    unsigned long long ra, rb, rd;
    asm("mov.b64 %0, {%1, %2};" : "=l"(ra) : "f"(a.x), "f"(a.y));
    asm("mov.b64 %0, {%1, %2};" : "=l"(rb) : "f"(b.x), "f"(b.y));
    asm("add.rn.f32x2 %0, %1, %2;" : "=l"(rd) : "l"(ra), "l"(rb));
    float2 d;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(rd));
    return d;
}

struct PairCursor {
    int i, m, m0, m1, tile, ord;      // unit slot i of this CTA, model m in [m0, m1), ordinal of the (non-empty) unit
    bool ok;
};

// ---- prep: one CTA per slot; owned slots get a dense index and an operand image ----------------
__global__ void __launch_bounds__(256) prep_images_f16_kernel(const float* __restrict__ slab, const int* __restrict__ slot_client,
                                                              int first_client, int n_local, int n_slots,
                                                              unsigned char* __restrict__ images, int* __restrict__ owned_slots,
                                                              int* __restrict__ n_owned) {
    __shared__ int s_idx;
    __shared__ float w2s[Z * H], w3s[H * Z], b2s[8], b3s[32];
    const int slot = blockIdx.x, tid = threadIdx.x;
    const int c = (slot_client ? slot_client[slot] : first_client + slot) - first_client;
    if (c < 0 || c >= n_local) return;
    if (tid == 0) { s_idx = atomicAdd(n_owned, 1); owned_slots[s_idx] = slot; }
    const float* row = slab + (size_t)c * P_PAD;
    for (int i = tid; i < Z * H; i += 256) { w2s[i] = row[OFF_W2 + i]; w3s[i] = row[OFF_W3 + i]; }
    if (tid < Z) b2s[tid] = row[OFF_B2 + tid];
    if (tid < H) b3s[tid] = row[OFF_B3 + tid];
    __syncthreads();
    unsigned char* img = images + (size_t)s_idx * FM_TC_IMAGE_BYTES;
    for (int i = tid; i < IMG_F_BYTES / 16; i += 256) reinterpret_cast<uint4*>(img)[i] = make_uint4(0, 0, 0, 0);
    __syncthreads();
    auto put = [&](int off_hi, int off_lo, int off, float v) {
        v = fminf(fmaxf(v, -F16_MAX), F16_MAX);
        const __half h = __float2half_rn(v);
        *reinterpret_cast<__half*>(img + off_hi + off) = h;
        *reinterpret_cast<__half*>(img + off_lo + off) = __float2half_rn(v - __half2float(h));
    };
    for (int p = tid; p < NP; p += 256) {
        const float v = row[p];
        if (p < OFF_B1) { const int n = p / D, k = p - n * D; put(IMG_W1, IMG_W1LO, (k >> 6) * 4096 + sw128h(n, k & 63), v); }
        else if (p < OFF_W2) put(IMG_W1, IMG_W1LO, 4096 + sw128h(p - OFF_B1, 115 & 63), v);     // b1 rides in column 115 (x[115] == 1)
        else if (p >= OFF_W4 && p < OFF_B4) { const int q = p - OFF_W4, d = q / H, j = q - d * H; put(IMG_W4, IMG_W4LO, sw64h(d, j), v); }
        else if (p >= OFF_B4) put(IMG_W4, IMG_W4LO, sw64h(p - OFF_B4, 27), v);                 // b4 rides in column 27 (h2[27] == 1)
    }
    if (tid == 0) {
        put(IMG_W1, IMG_W1LO, 4096 + sw128h(27, 115 & 63), 1.f);      // h1[27] = relu(1 * x[115]) = 1
        put(IMG_W32, IMG_W32LO, sw64h(27, 27), 1.f);                  // h2[27] = relu(1 * h1[27]) = 1
    }
    // W32 = W3 W2 (27 x 27), b32 = W3 b2 + b3 in column 27
    for (int e = tid; e < H * (H + 1); e += 256) {
        const int n = e / (H + 1), k = e - n * (H + 1);
        float acc = k < H ? 0.f : b3s[n];
#pragma unroll
        for (int i = 0; i < Z; ++i) acc = fmaf(w3s[n * Z + i], k < H ? w2s[i * H + k] : b2s[i], acc);
        put(IMG_W32, IMG_W32LO, sw64h(n, k), acc);
    }
}

// TIMING: CTA 0 accumulates SM cycles per wait / work item into dbg[0..23] (see the TK_* slots), for tuning only
template <bool TIMING>
__global__ void __launch_bounds__(NT_F, 1)
score_fused_f16_kernel(const __grid_constant__ CUtensorMap xmap, const unsigned char* __restrict__ images,
                       const int* __restrict__ owned_slots, const int* __restrict__ n_owned_ptr, int x_rows, int n_chunks,
                       float* __restrict__ partial, float* __restrict__ out_mse, unsigned* __restrict__ gbar,
                       unsigned long long* __restrict__ dbg) {
    extern __shared__ __align__(1024) unsigned char smem_dyn[];
    FSmem& S = *reinterpret_cast<FSmem*>(smem_dyn);
    if ((smem_u32(smem_dyn) & 1023u) != 0u) __trap();         // 128-B swizzle needs a 1024-B aligned base
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int n_models = *n_owned_ptr;
    if (n_models <= 0) return;                                // uniform over the grid
    const int n_tiles = (x_rows + TROWS - 1) / TROWS;
    const int n_units = n_tiles * n_chunks;
    const int my_units = (int)blockIdx.x < n_units ? (n_units - (int)blockIdx.x + (int)gridDim.x - 1) / (int)gridDim.x : 0;

    auto next_unit = [&](PairCursor& c) {
        for (;;) {
            if (++c.i >= my_units) { c.ok = false; return; }
            const int u = (int)blockIdx.x + c.i * (int)gridDim.x, ch = u % n_chunks;
            c.tile = u / n_chunks;
            c.m = c.m0 = ch * n_models / n_chunks;          // n_models <= FM_MAX_SEL, n_chunks <= 4: no overflow
            c.m1 = (ch + 1) * n_models / n_chunks;
            if (c.m < c.m1) { ++c.ord; c.ok = true; return; }
        }
    };
    auto first_pair = [&]() { PairCursor c; c.i = -1; c.m = c.m0 = c.m1 = 0; c.tile = 0; c.ord = -1; c.ok = false; next_unit(c); return c; };
    auto next_pair = [&](PairCursor& c) { if (++c.m >= c.m1) next_unit(c); };

    if (tid == 0) {
        mbar_init(&S.bar_x, 1); mbar_init(&S.bar_xfree, 384); mbar_init(&S.bar_xt, 128);
        for (int b = 0; b < NW; ++b) {
            mbar_init(&S.bar_w1[b], 1); mbar_init(&S.bar_w32[b], 1); mbar_init(&S.bar_w4[b], 1);
        }
        for (int b = 0; b < 2; ++b) mbar_init(&S.bar_d2free[b], 256);
        for (int b = 0; b < RB; ++b) { mbar_init(&S.bar_m1[b], 1); mbar_init(&S.bar_m1b[b], 1); mbar_init(&S.bar_m2[b], 1); }
        for (int b = 0; b < NB; ++b) {
            mbar_init(&S.bar_h1[b], 128); mbar_init(&S.bar_h2[b], 128);
        }
        fence_mbar_init();
    }
    if (warp == MMA1_WARP) {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(F_TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    for (int i = tid; i < n_models * 256; i += NT_F) partial[(size_t)blockIdx.x * n_models * 256 + i] = 0.f;
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    constexpr uint32_t IDESC_N32 = idesc_f16(128, 32), IDESC_N128 = idesc_f16(128, 128);

    // total pairs of this CTA and, per role, a walk over them in the same order
    const bool has_work = first_pair().ok;
    if (!has_work) {
    } else if (warp == MMA1_WARP || warp == MMA1B_WARP || warp == MMA2_WARP) {
        // Three issuing threads, one per GEMM stage.  ncu of the single-issuer versions showed the tensor pipe active
        // for exactly its ~0.93 k cycles of math per pair and idle for the ~0.8-1.0 k cycles the issuing thread spent
        // on everything else (waits, phase arithmetic, descriptor selection): the MMA queue drains while its only
        // producer does bookkeeping, so period = math + issue overhead, whatever the buffering.  With one issuer per
        // stage the bookkeeping of one overlaps the MMAs of the others.  Each thread walks the pairs in order with
        // blocking waits; every cross-thread hazard is covered by a commit barrier ("these MMAs have RETIRED"):
        //   MMA1(q)  overwrites D1[q&1]   -> waits bar_m1b of pair q-2 (MMA1b(q-2) retired)
        //   MMA1b(q) overwrites D1b[q&1]  -> waits bar_m2 of pair q-2  (MMA2(q-2) retired)
        //   MMA2(q)  overwrites D2[q&1]   -> waits bar_d2free  (epilogue 2 of pair q-2 has read it)
        // No thread ever waits for something that needs its own later work, so units of any length are safe.
        if (elect_one()) {
            int n_pairs = 0;
            { PairCursor t = first_pair(); while (t.ok) { ++n_pairs; next_pair(t); } }
            int s = 0;                 // image stage q % NW and its phase (q / NW) & 1, kept incrementally
            uint32_t sph = 0;
            const bool tk_on = TIMING && blockIdx.x == 0;
            long long tk[TK_N] = {};
            long long tk_t = TIMING ? clock64() : 0;
            const long long tk_0 = tk_t;
            if (warp == MMA1_WARP) {
                const uint64_t dw1 = smem_desc_sw128(S.w1[0]);
                PairCursor c1 = first_pair();
                int unit1 = -1;
                for (int q = 0; q < n_pairs; ++q) {
                    const int r = q & 1;
                    if (c1.ord != unit1) {
                        if (tk_on && c1.ord >= 1 && c1.ord <= 4) dbg[32 + 8 * (c1.ord - 1)] = (unsigned long long)clock64();
                        mbar_wait(&S.bar_xt, (uint32_t)(c1.ord & 1));                                     // tile parked
                        if (tk_on && c1.ord >= 1 && c1.ord <= 4) dbg[33 + 8 * (c1.ord - 1)] = (unsigned long long)clock64();
                        unit1 = c1.ord;
                    }
                    FT(TK_1_XT);
                    mbar_wait(&S.bar_w1[s], sph);
                    FT(TK_1_W);
                    if (q >= NB) mbar_wait(&S.bar_m1b[(q - NB) % RB], (uint32_t)(((q - NB) / RB) & 1));
                    FT(TK_1_FREE);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 0 * TRACE_N + (q)] = (unsigned long long)clock64();
                    tc_fence_after();
                    const uint32_t d1 = tmem + C_D1 + 32 * r, xh = tmem + C_X, xl = tmem + C_XLO;
                    const uint64_t bw = dw1 + (uint64_t)((s * 2 * W1_BYTES) >> 4);
#pragma unroll
                    for (int kk = 0; kk < 8; ++kk) {
                        const uint64_t bh = bw + (uint64_t)(((kk >> 2) * 4096 + (kk & 3) * 32) >> 4);
                        const uint64_t bl = bh + (uint64_t)(W1_BYTES >> 4);
                        mma_ts_f16(d1, xh + kk * 8, bh, IDESC_N32, kk ? 1u : 0u);
                        mma_ts_f16(d1, xh + kk * 8, bl, IDESC_N32, 1u);
                        mma_ts_f16(d1, xl + kk * 8, bh, IDESC_N32, 1u);
                    }
                    tc_commit(&S.bar_m1[q % RB]);
                    next_pair(c1);
                    if (++s == NW) { s = 0; sph ^= 1u; }
                    FT(TK_1_ISSUE);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 1 * TRACE_N + (q)] = (unsigned long long)clock64();
                }
                if (tk_on) tk[TK_TOTAL] = clock64() - tk_0;
            } else if (warp == MMA1B_WARP) {
                const uint64_t dw32 = smem_desc_sw64(S.w32[0]);
                for (int q = 0; q < n_pairs; ++q) {
                    const int r = q & 1;
                    mbar_wait(&S.bar_h1[r], (uint32_t)((q >> 1) & 1));
                    FT(TK_1B_H);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 4 * TRACE_N + (q)] = (unsigned long long)clock64();
                    mbar_wait(&S.bar_w32[s], sph);
                    FT(TK_1B_W);
                    if (q >= NB) mbar_wait(&S.bar_m2[(q - NB) % RB], (uint32_t)(((q - NB) / RB) & 1));
                    FT(TK_1B_FREE);
                    tc_fence_after();
                    const uint32_t d = tmem + C_D1B + 32 * r, ah = tmem + C_D1 + 32 * r;
                    const uint64_t bw = dw32 + (uint64_t)((s * 2 * W32_BYTES) >> 4);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const uint64_t bh = bw + (uint64_t)((ks * 32) >> 4), bl = bh + (uint64_t)(W32_BYTES >> 4);
                        mma_ts_f16(d, ah + ks * 8, bh, IDESC_N32, ks ? 1u : 0u);
                        mma_ts_f16(d, ah + ks * 8, bl, IDESC_N32, 1u);
                    }
                    tc_commit(&S.bar_m1b[q % RB]);
                    if (++s == NW) { s = 0; sph ^= 1u; }
                    FT(TK_1B_ISSUE);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 5 * TRACE_N + (q)] = (unsigned long long)clock64();
                }
            } else {
                const uint64_t dw4 = smem_desc_sw64(S.w4[0]);
                for (int q = 0; q < n_pairs; ++q) {
                    const int r = q & 1;
                    mbar_wait(&S.bar_h2[r], (uint32_t)((q >> 1) & 1));
                    FT(TK_2_H);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 8 * TRACE_N + (q)] = (unsigned long long)clock64();
                    mbar_wait(&S.bar_w4[s], sph);
                    FT(TK_2_W);
                    if (q >= 2) mbar_wait(&S.bar_d2free[r], (uint32_t)(((q - 2) >> 1) & 1));
                    FT(TK_2_FREE);
                    tc_fence_after();
                    const uint32_t d = tmem + C_D2 + 128 * r, ah = tmem + C_D1B + 32 * r;
                    const uint64_t bw = dw4 + (uint64_t)((s * 2 * W4_BYTES) >> 4);
#pragma unroll
                    for (int ks = 0; ks < 2; ++ks) {
                        const uint64_t bh = bw + (uint64_t)((ks * 32) >> 4), bl = bh + (uint64_t)(W4_BYTES >> 4);
                        mma_ts_f16(d, ah + ks * 8, bh, IDESC_N128, ks ? 1u : 0u);
                        mma_ts_f16(d, ah + ks * 8, bl, IDESC_N128, 1u);
                    }
                    tc_commit(&S.bar_m2[q % RB]);
                    if (++s == NW) { s = 0; sph ^= 1u; }
                    FT(TK_2_ISSUE);
                    if (TIMING && tk_on && (q) < TRACE_N) dbg[TRACE_BASE + 9 * TRACE_N + (q)] = (unsigned long long)clock64();
                }
            }
            if (tk_on) for (int i = 0; i < TK_N; ++i) if (tk[i]) dbg[i] = (unsigned long long)tk[i];
        }
        __syncwarp();
    } else if (warp == LOAD_WARP) {
        // One thread feeds the shared memory: weight images NW pairs ahead (image l of a kind goes into stage l % NW
        // once the MMAs that read image l - NW have retired: the group's completion barrier bar_m*[(l - NW) % RB]; the
        // same slot's next completion, RB pairs later, needs images this thread requests only after the test succeeded,
        // so no phase can be missed) and the fp32 tile of the next unit (once the parking warps and the squared-error warps both have
        // their copy of the current one).  Everything is polled, so no kind of load can hold back another.
        if (elect_one()) {
            int n_pairs = 0, n_units_ne = 0;
            { PairCursor t = first_pair(); while (t.ok) { ++n_pairs; n_units_ne = t.ord + 1; next_pair(t); } }
            PairCursor cw1 = first_pair(), cw32 = cw1, cw4 = cw1, cx = cw1;
            int l1 = 0, l32 = 0, l4 = 0, lx = 0;
            auto load_img = [&](unsigned char* dst, int m, int off, uint32_t bytes, uint64_t* bar) {
                mbar_expect_tx(bar, bytes);
                bulk_load(dst, images + (size_t)m * FM_TC_IMAGE_BYTES + off, bytes, bar);
            };
            while (l1 < n_pairs || l32 < n_pairs || l4 < n_pairs || lx < n_units_ne) {
                if (lx < n_units_ne && (lx == 0 || mbar_test(&S.bar_xfree, (uint32_t)((lx - 1) & 1)))) {
                    mbar_expect_tx(&S.bar_x, 4u * XBLK);
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb) tma_load_2d(S.x + kb * XBLK, &xmap, kb * 32, cx.tile * TROWS, &S.bar_x);
                    if (TIMING && blockIdx.x == 0 && lx >= 1 && lx <= 4) dbg[37 + 8 * (lx - 1)] = (unsigned long long)clock64();
                    next_unit(cx); ++lx;
                }
                if (l1 < n_pairs && (l1 < NW || mbar_test(&S.bar_m1[(l1 - NW) % RB], (uint32_t)(((l1 - NW) / RB) & 1)))) {
                    load_img(S.w1[l1 % NW], cw1.m, IMG_W1, 2 * W1_BYTES, &S.bar_w1[l1 % NW]); next_pair(cw1); ++l1;
                }
                if (l32 < n_pairs && (l32 < NW || mbar_test(&S.bar_m1b[(l32 - NW) % RB], (uint32_t)(((l32 - NW) / RB) & 1)))) {
                    load_img(S.w32[l32 % NW], cw32.m, IMG_W32, 2 * W32_BYTES, &S.bar_w32[l32 % NW]); next_pair(cw32); ++l32;
                }
                if (l4 < n_pairs && (l4 < NW || mbar_test(&S.bar_m2[(l4 - NW) % RB], (uint32_t)(((l4 - NW) / RB) & 1)))) {
                    load_img(S.w4[l4 % NW], cw4.m, IMG_W4, 2 * W4_BYTES, &S.bar_w4[l4 % NW]); next_pair(cw4); ++l4;
                }
            }
        }
        __syncwarp();
    } else if (warp < 4) {
        // ---- group A: parks tiles in TMEM, epilogues 1a (h1) and 1b (h2) of every pair ----
        const int r = warp * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
        uint32_t v[32], hi[16], lo[16];     // lo: tile parking only
        auto park_x = [&](int ut) {                   // fp32 tile -> fp16 head + remainder, one row per thread, k = 115 := 1
            mbar_wait(&S.bar_x, (uint32_t)(ut & 1));
            const uint32_t xb_t = lane_base;
#pragma unroll
            for (int cb = 0; cb < 4; ++cb) {
                const unsigned char* xb = S.x + cb * XBLK + r * 128;
#pragma unroll
                for (int q4 = 0; q4 < 8; ++q4) {
                    float4 xv = *reinterpret_cast<const float4*>(xb + ((q4 ^ (r & 7)) << 4));
                    if (cb == 3 && q4 == 4) xv.w = 1.f;                    // k = 96 + 16 + 3 = 115: the bias column
                    split2(xv.x, xv.y, hi[q4 * 2], lo[q4 * 2]);
                    split2(xv.z, xv.w, hi[q4 * 2 + 1], lo[q4 * 2 + 1]);
                }
                tmem_st16(xb_t + C_X + cb * 16, hi);
                tmem_st16(xb_t + C_XLO + cb * 16, lo);
            }
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(&S.bar_xt);
            mbar_arrive(&S.bar_xfree);
        };
        auto relu_split = [&](uint32_t col, uint64_t* bar_in, uint32_t ph, uint64_t* bar_out) {
            mbar_wait(bar_in, ph);
            tc_fence_after();
            tmem_ld32(lane_base + col, v);
            // activations travel as ONE fp16 (ReLU + saturation + packing in a single convert per pair of values): their
            // rounding (2^-11, unbiased, different for every row) averages out over the 10^5..10^6 rows of an MSE, unlike a
            // rounding of the weights, which is the same for every row -- the weights keep head + remainder
#pragma unroll
            for (int j = 0; j < 16; ++j) hi[j] = pack_f16x2_relu_sat(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1]));
            tmem_st16(lane_base + col, hi);            // in place: the fp16 operand replaces the fp32 accumulator it came from
            tmem_st_wait();
            tc_fence_before();
            mbar_arrive(bar_out);
        };
        auto e1 = [&](int q) { const int rb = q & (NB - 1); relu_split(C_D1 + 32 * rb, &S.bar_m1[q % RB], (uint32_t)((q / RB) & 1), &S.bar_h1[rb]); };
        auto e1b = [&](int q) { const int rb = q & (NB - 1); relu_split(C_D1B + 32 * rb, &S.bar_m1b[q % RB], (uint32_t)((q / RB) & 1), &S.bar_h2[rb]); };
        int n_pairs = 0, n_units_ne = 0;
        { PairCursor t = first_pair(); while (t.ok) { ++n_pairs; n_units_ne = t.ord + 1; next_pair(t); } }
        // Three kinds of work, each taken in order as soon as its input is there (warp-uniform non-blocking tests):
        //   e1(qe1)    when MMA1(qe1) has retired,   e1b(qe1b)  when MMA1b(qe1b) has retired,
        //   park(unit) when the unit's fp32 tile has landed in the staging buffer AND the X columns of TMEM are free:
        //              their last reader was MMA1 of the last pair of unit - 1, whose retirement this warp has seen once
        //              the next e1 pair belongs to this unit (or there is none left).
        // A fixed order would deadlock on units shorter than three pairs (the next tile only lands after the
        // squared-error warps entered the current unit, which needs e1b of its first pair).
        PairCursor ce1 = first_pair();                // pair e1 handles next
        int qe1 = 0, qe1b = 0, parked = 0;
        auto uniform = [&](bool ok) -> bool { return __all_sync(0xffffffffu, ok) != 0; };
        const bool tk_on = TIMING && blockIdx.x == 0 && tid == 0;
        long long tk[TK_N] = {};
        long long tk_t = TIMING ? clock64() : 0;
        while (qe1b < n_pairs) {
            bool did = false;
            FT(TK_A_IDLE);
            if (parked < n_units_ne && (qe1 >= n_pairs || ce1.ord >= parked) &&
                uniform(mbar_test(&S.bar_x, (uint32_t)(parked & 1)))) {
                if (tk_on && parked >= 1 && parked <= 4) dbg[34 + 8 * (parked - 1)] = (unsigned long long)clock64();
                park_x(parked);
                if (tk_on && parked >= 1 && parked <= 4) dbg[35 + 8 * (parked - 1)] = (unsigned long long)clock64();
                ++parked;
                did = true;
                FT(TK_A_PARK);
            }
            if (qe1 < n_pairs && uniform(mbar_test(&S.bar_m1[qe1 % RB], (uint32_t)((qe1 / RB) & 1)))) {
                if (TIMING && tk_on && (qe1) < TRACE_N) dbg[TRACE_BASE + 2 * TRACE_N + (qe1)] = (unsigned long long)clock64();
                e1(qe1);
                { const int o0 = ce1.ord; next_pair(ce1);
                  if (tk_on && ce1.ord != o0 && o0 < 4) dbg[36 + 8 * o0] = (unsigned long long)clock64(); }   // last e1 of unit o0 done
                ++qe1;
                did = true;
                FT(TK_A_E1);
                if (TIMING && tk_on && (qe1 - 1) < TRACE_N) dbg[TRACE_BASE + 3 * TRACE_N + (qe1 - 1)] = (unsigned long long)clock64();
            }
            if (qe1b < qe1 && uniform(mbar_test(&S.bar_m1b[qe1b % RB], (uint32_t)((qe1b / RB) & 1)))) {
                if (TIMING && tk_on && (qe1b) < TRACE_N) dbg[TRACE_BASE + 6 * TRACE_N + (qe1b)] = (unsigned long long)clock64();
                e1b(qe1b);
                ++qe1b;
                did = true;
                FT(TK_A_E1B);
                if (TIMING && tk_on && (qe1b - 1) < TRACE_N) dbg[TRACE_BASE + 7 * TRACE_N + (qe1b - 1)] = (unsigned long long)clock64();
            }
            if (!did) __nanosleep(20);
        }
        if (tk_on) for (int i = TK_A_PARK; i <= TK_A_IDLE; ++i) dbg[i] = (unsigned long long)tk[i];
    } else {
        // ---- group B: epilogue 2 (squared error against the row kept in registers), partial sums.  Eight warps: warps
        //      4-7 take output columns 0..63 of their lane quarter, warps 8-11 columns 64..127, so a pair costs each
        //      thread two 32-column loads and 32 packed subtract / multiply-adds ----
        const int wq = warp & 3, half = (warp - 4) >> 2;
        const int r = wq * 32 + lane;
        const uint32_t lane_base = tmem + ((uint32_t)(wq * 32) << 16);
        uint32_t v[32];
        float2 nx[32];                              // -x of this thread's half row (64 floats), for the whole unit
        // per-thread running sums lanes[cta][model][256]: every slot has exactly one writer, in program order, so the
        // fire-and-forget reductions below are deterministic and cost no shuffle, no wait and no address arithmetic
        float* my_lane = partial + ((size_t)blockIdx.x * n_models) * 256 + (warp - 4) * 32 + lane;
        PairCursor c = first_pair();
        int g = 0, ut = 0;
        bool fresh = true;
        const bool tk_on = TIMING && blockIdx.x == 0 && tid == 128;
        long long tk[TK_N] = {};
        long long tk_t = TIMING ? clock64() : 0;
        while (c.ok) {
            FT(TK_B_REST);
            if (fresh) {
                mbar_wait(&S.bar_x, (uint32_t)(ut & 1));
#pragma unroll
                for (int q = 0; q < 16; ++q) {
                    const int qq = half * 16 + q;     // float4 index inside the 128-float row
                    const float4 xv = *reinterpret_cast<const float4*>(S.x + (qq >> 3) * XBLK + r * 128 + (((qq & 7) ^ (r & 7)) << 4));
                    // columns 115..127 of D2 are exact zeros (no such rows in W4): their x is padding
                    nx[q * 2] = make_float2(qq * 4 < D ? -xv.x : 0.f, qq * 4 + 1 < D ? -xv.y : 0.f);
                    nx[q * 2 + 1] = make_float2(qq * 4 + 2 < D ? -xv.z : 0.f, qq * 4 + 3 < D ? -xv.w : 0.f);
                }
                mbar_arrive(&S.bar_xfree);
                fresh = false;
                FT(TK_B_ENTRY);
            }
            const int b = g & 1;
            mbar_wait(&S.bar_m2[g % RB], (uint32_t)((g / RB) & 1));
            tc_fence_after();
            FT(TK_B_WAIT);
            if (TIMING && tk_on && (g) < TRACE_N) dbg[TRACE_BASE + 10 * TRACE_N + (g)] = (unsigned long long)clock64();
            float2 acc[4] = {make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f), make_float2(0.f, 0.f)};
            auto acc_block = [&](int cb) {            // four independent chains: the FMAs pipeline
#pragma unroll
                for (int j = 0; j < 16; ++j) {
                    const float2 e = fadd2(make_float2(__uint_as_float(v[2 * j]), __uint_as_float(v[2 * j + 1])), nx[cb * 16 + j]);
                    acc[j & 3] = ffma2(e, e, acc[j & 3]);
                }
            };
            const uint32_t d2 = lane_base + C_D2 + 128 * b + 64 * half;
            tmem_ld32(d2, v);
            acc_block(0);
            tmem_ld32(d2 + 32, v);
            tc_fence_before();
            mbar_arrive(&S.bar_d2free[b]);            // this thread's share of D2[b] is in registers
            FT(TK_B_HOLD);
            if (TIMING && tk_on && (g) < TRACE_N) dbg[TRACE_BASE + 11 * TRACE_N + (g)] = (unsigned long long)clock64();
            acc_block(1);
            const int row = c.tile * TROWS + r;
            const float sse = row < x_rows ? ((acc[0].x + acc[0].y) + (acc[1].x + acc[1].y)) + ((acc[2].x + acc[2].y) + (acc[3].x + acc[3].y)) : 0.f;
            atomicAdd(my_lane + (size_t)c.m * 256, sse);              // result unused: a RED; nobody else writes this slot
            if (TIMING && tk_on && (g) < TRACE_N) dbg[TRACE_BASE + 12 * TRACE_N + (g)] = (unsigned long long)clock64();
            PairCursor cn = c;
            next_pair(cn);
            if (!cn.ok || cn.i != c.i) { ++ut; fresh = true; }
            c = cn;
            ++g;
        }
        if (tk_on) for (int i = TK_B_WAIT; i <= TK_B_ENTRY; ++i) dbg[i] = (unsigned long long)tk[i];
    }
    __threadfence();                         // this thread's running sums are visible device-wide before the grid barrier
    tc_fence_before();
    __syncthreads();
    if (warp == MMA1_WARP) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(F_TMEM_COLS));

    // ---- grid-wide barrier (the grid never exceeds the SM count and a CTA fills an SM: all CTAs are resident), then
    //      the per-model MSE from the partial sums in a fixed order ----
    if (tid == 0) {
        __threadfence();
        const unsigned gen = *reinterpret_cast<volatile unsigned*>(gbar + 1);
        const unsigned old = atomicAdd(gbar, 1u);
        if (old == gridDim.x - 1) {
            *reinterpret_cast<volatile unsigned*>(gbar) = 0u;
            __threadfence();
            atomicAdd(gbar + 1, 1u);
        } else {
            while (*reinterpret_cast<volatile unsigned*>(gbar + 1) == gen) __nanosleep(32);
        }
        __threadfence();
    }
    __syncthreads();
    for (int m = blockIdx.x; m < n_models; m += gridDim.x) {
        double s = 0.0;
        for (int i = tid; i < (int)gridDim.x * 256; i += NT_F)       // fixed order: thread-strided, then warps in order
            s += (double)__ldcg(partial + ((size_t)(i >> 8) * n_models + m) * 256 + (i & 255));
        s = warp_sum(s);
        if (lane == 0) S.red[warp] = s;
        __syncthreads();
        if (tid == 0) {
            double t = 0.0;
            for (int w = 0; w < NT_F / 32; ++w) t += S.red[w];
            out_mse[owned_slots[m]] = (float)(t / ((double)x_rows * (double)D));
        }
        __syncthreads();
    }
    (void)dbg;
}

}  // namespace tcf
}  // namespace fm

namespace {
typedef CUresult (*EncodeTiledFn3)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                   const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                   CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
int encode_x_map_f(CUtensorMap* map, const float* base, int64_t rows) {
    static EncodeTiledFn3 fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) != cudaSuccess ||
            q != cudaDriverEntryPointSuccess)
            return (int)cudaErrorNotSupported;
        fn = reinterpret_cast<EncodeTiledFn3>(p);
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

extern "C" int64_t fm_score_fused_scratch_floats(int64_t x_rows, int32_t n_models) {
    const int64_t n_tiles = (x_rows + fm::tc::TROWS - 1) / fm::tc::TROWS;
    (void)n_tiles;                                   // per-thread running sums: [CTAs <= 160][models][256]
    return 160 * 256 * (int64_t)(n_models > 0 ? n_models : 1);
}

extern "C" int fm_score_fused_f16(const FmFusedScoreArgs* a, void* stream) {
    using namespace fm;
    using namespace fm::tc;
    using namespace fm::tcf;
    if (a->n_slots <= 0 || a->x_rows <= 0) return 0;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t err = cudaMemsetAsync(a->n_owned, 0, sizeof(int), st);
    if (err != cudaSuccess) return (int)err;
    prep_images_f16_kernel<<<a->n_slots, 256, 0, st>>>(a->slab, a->slot_client, a->first_client, a->n_local, a->n_slots,
                                                       a->images, a->owned_slots, a->n_owned);
    CUtensorMap map;
    int rc = encode_x_map_f(&map, a->X, a->x_rows);
    if (rc) return rc;
    static int sms = 0;
    if (!sms) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const size_t smem = sizeof(FSmem);
    auto kern = a->dbg_cycles ? score_fused_f16_kernel<true> : score_fused_f16_kernel<false>;
    err = cudaFuncSetAttribute((const void*)kern, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    // chunks of models per tile: balance the units over the SMs (same cost model as score_multi_tc.cu)
    const int n_tiles = (int)((a->x_rows + TROWS - 1) / TROWS);
    const int m_est = a->n_slots < a->n_local ? a->n_slots : (a->n_local > 0 ? a->n_local : 1);
    int n_chunks = 1;
    double best = 1e30;
    for (int c = 1; c <= 4 && c <= (m_est > 0 ? m_est : 1); ++c) {
        const long long units = (long long)n_tiles * c;
        const long long g = units < sms ? units : sms;
        const double cost = (double)((units + g - 1) / g) * ((double)m_est / c + 4.0);
        if (cost < best * 0.98) { best = cost; n_chunks = c; }
    }
    const long long units = (long long)n_tiles * n_chunks;
    const int grid = (int)(units < sms ? units : sms);
    kern<<<grid, NT_F, smem, st>>>(map, a->images, a->owned_slots, a->n_owned, (int)a->x_rows, n_chunks, a->partial,
                                   a->out_mse, a->grid_barrier, a->dbg_cycles);
    return (int)cudaGetLastError();
}
