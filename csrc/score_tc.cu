// Row-scoring forward on the 5th-generation tensor cores (tcgen05 + TMEM + TMA), TF32 inputs,
// fp32 accumulation.  Same contract as score_rows_kernel (score_kernel.cu) with 128-row tiles.
//
// This is the GEMM-shaped work of the framework: one model applied to 10^4..10^6 rows (MSE-weight
// scoring on the pooled dev set, client_trainer.py:115-130; evaluation, evaluator.py:56-94; vote and
// verification forwards).  Per 128-row tile:
//
//   TMA   : X tile [128 x 128] fp32 -> smem, four 128-B-swizzled K-blocks (cp.async.bulk.tensor.2d)
//   MMA 1 : D1[128 x 32]  = X[128 x 128] . W1^T            16 x tcgen05.mma kind::tf32 (SS), M=128 N=32 K=8
//   epi 1 : one thread per row: tcgen05.ld D1 -> +b1, ReLU -> 27->7->27 bottleneck in registers (FFMA,
//           weights broadcast from smem) -> latent out; h2 -> tcgen05.st into TMEM as the next A operand
//   MMA 2 : D2[128 x 128] = H2[128 x 32] . W4^T            4 x tcgen05.mma kind::tf32 (TS: A from TMEM)
//   epi 2 : tcgen05.ld D2 -> +b4, minus x (swizzled smem reads, conflict-free), squared error -> out
//
// Warp roles: warps 0-3 own TMEM lanes 32w..32w+31 (= tile rows) and run both epilogues; one elected
// lane of warp 4 issues TMA and MMA; completion travels through mbarriers (tcgen05.commit).  Two CTAs
// per SM (256 TMEM columns each) overlap one tile's epilogue with the other's copy and MMAs.
#include "tc_common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {
namespace tc {

constexpr int NT_TC = 160;                 // 4 epilogue warps + 1 producer/MMA warp
constexpr int TMEM_COLS = 256;
constexpr int COL_D1A = 0, COL_D1B = 32, COL_A2 = 64, COL_A2LO = 96, COL_D2 = 128;

// PREC = 1: single TF32 pass.  PREC = 3: error-compensated 3xTF32 -- every operand is split into
// hi = value truncated to TF32 and lo = value - hi (exact in fp32), and each product is formed as
// hi*hi + hi*lo + lo*hi with fp32 accumulation, which recovers ~22 mantissa bits (fp32-grade results).
template <int PREC>
struct TcSmem {
    static constexpr int LO = PREC == 3 ? 1 : 0;
    alignas(1024) unsigned char x[4 * XBLK];      // 64 KB, 4 K-blocks, SWIZZLE_128B (PREC 3: hi part, in place)
    alignas(1024) unsigned char w1[4 * W1BLK];    // 16 KB, B operand of MMA 1: [N=32][K=128] K-major
    alignas(1024) unsigned char w4[128 * 128];    // 16 KB, B operand of MMA 2: [N=128][K=32] K-major
    alignas(1024) unsigned char xlo[LO ? 4 * XBLK : 1024];
    alignas(1024) unsigned char w1lo[LO ? 4 * W1BLK : 1024];
    alignas(1024) unsigned char w4lo[LO ? 128 * 128 : 1024];
    alignas(16) float w2[Z * 28];                 // [i][j]
    alignas(16) float w3[H * 8];                  // [j][i]
    alignas(16) float b1[32], b2[8], b3[32];
    alignas(8) uint64_t bar_x, bar_d1[2], bar_h2, bar_d2;
    uint32_t tmem_base;
};

// stage one model into the UMMA operand layouts (TF32, round-to-nearest) and plain fp32 tables
template <int PREC>
__device__ __forceinline__ void stage_model(TcSmem<PREC>& S, const float* __restrict__ row) {
    const int tid = threadIdx.x;
    for (int i = tid; i < (4 * W1BLK) / 16; i += NT_TC) reinterpret_cast<uint4*>(S.w1)[i] = make_uint4(0, 0, 0, 0);
    for (int i = tid; i < (128 * 128) / 16; i += NT_TC) reinterpret_cast<uint4*>(S.w4)[i] = make_uint4(0, 0, 0, 0);
    if (PREC == 3) {
        for (int i = tid; i < (4 * W1BLK) / 16; i += NT_TC) reinterpret_cast<uint4*>(S.w1lo)[i] = make_uint4(0, 0, 0, 0);
        for (int i = tid; i < (128 * 128) / 16; i += NT_TC) reinterpret_cast<uint4*>(S.w4lo)[i] = make_uint4(0, 0, 0, 0);
    }
    for (int i = tid; i < Z * 28; i += NT_TC) S.w2[i] = 0.f;
    for (int i = tid; i < H * 8; i += NT_TC) S.w3[i] = 0.f;
    if (tid < 32) { S.b1[tid] = 0.f; S.b3[tid] = 0.f; }
    if (tid < 8) S.b2[tid] = 0.f;
    __syncthreads();
    for (int p = tid; p < NP; p += NT_TC) {
        const float v = row[p];
        if (p < OFF_B1) {
            const int n = p / D, k = p - n * D;
            const int off = (k >> 5) * W1BLK + sw128(n, k & 31);
            if (PREC == 3) {
                const uint32_t hi = trunc_tf32(v);
                *reinterpret_cast<uint32_t*>(S.w1 + off) = hi;
                *reinterpret_cast<float*>(S.w1lo + off) = v - __uint_as_float(hi);
            } else *reinterpret_cast<uint32_t*>(S.w1 + off) = cvt_tf32(v);
        } else if (p < OFF_W2) S.b1[p - OFF_B1] = v;
        else if (p < OFF_B2) { const int q = p - OFF_W2, i = q / H, j = q - i * H; S.w2[i * 28 + j] = v; }
        else if (p < OFF_W3) S.b2[p - OFF_B2] = v;
        else if (p < OFF_B3) { const int q = p - OFF_W3, j = q / Z, i = q - j * Z; S.w3[j * 8 + i] = v; }
        else if (p < OFF_W4) S.b3[p - OFF_B3] = v;
        else if (p < OFF_B4) {
            const int q = p - OFF_W4, d = q / H, j = q - d * H;
            const int off = sw128(d, j);
            if (PREC == 3) {
                const uint32_t hi = trunc_tf32(v);
                *reinterpret_cast<uint32_t*>(S.w4 + off) = hi;
                *reinterpret_cast<float*>(S.w4lo + off) = v - __uint_as_float(hi);
            } else *reinterpret_cast<uint32_t*>(S.w4 + off) = cvt_tf32(v);
        } else {                                    // b4[d] rides in column 27 of the W4 operand (h2[27] == 1)
            const int off = sw128(p - OFF_B4, 27);
            if (PREC == 3) {
                const uint32_t hi = trunc_tf32(v);
                *reinterpret_cast<uint32_t*>(S.w4 + off) = hi;
                *reinterpret_cast<float*>(S.w4lo + off) = v - __uint_as_float(hi);
            } else *reinterpret_cast<uint32_t*>(S.w4 + off) = cvt_tf32(v);
        }
    }
    fence_proxy_async();          // generic-proxy writes -> visible to the tensor core (async proxy)
    __syncthreads();
}

// residual shard: xlo = x - trunc_tf32(x), so that x == hi + lo exactly with hi = what the tensor core reads
__global__ void split_residual_kernel(const float4* __restrict__ x, float4* __restrict__ lo, size_t n4) {
    for (size_t i = blockIdx.x * (size_t)blockDim.x + threadIdx.x; i < n4; i += (size_t)gridDim.x * blockDim.x) {
        const float4 v = x[i];
        lo[i] = make_float4(v.x - __uint_as_float(trunc_tf32(v.x)), v.y - __uint_as_float(trunc_tf32(v.y)),
                            v.z - __uint_as_float(trunc_tf32(v.z)), v.w - __uint_as_float(trunc_tf32(v.w)));
    }
}

template <int PREC>
__global__ void __launch_bounds__(NT_TC, PREC == 3 ? 1 : 2)
score_rows_tc_kernel(const FmScoreArgs a, const __grid_constant__ CUtensorMap xmap,
                     const __grid_constant__ CUtensorMap xlomap, long long x_rows, int tiles_per_cta) {
    extern __shared__ unsigned char smem_dyn[];
    using SM = TcSmem<PREC>;
    SM& S = *reinterpret_cast<SM*>((reinterpret_cast<uintptr_t>(smem_dyn) + 1023) & ~(uintptr_t)1023);
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    // work units of `tiles_per_cta` consecutive tiles are dealt round-robin to the CTAs: with several ranks
    // only 1/world of the slots is owned locally, and interleaving keeps the owned work balanced
    const int unit = tiles_per_cta;
    if (blockIdx.x * unit >= a.n_tiles) return;
    const bool need_sse = a.out_sse != nullptr;

    if (tid == 0) {
        mbar_init(&S.bar_x, 1); mbar_init(&S.bar_d1[0], 1); mbar_init(&S.bar_d1[1], 1); mbar_init(&S.bar_d2, 1);
        mbar_init(&S.bar_h2, 128);
        fence_mbar_init();
    }
    if (warp == 4) {      // one warp allocates the CTA's tensor memory
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;" ::"r"(smem_u32(&S.tmem_base)), "r"(TMEM_COLS));
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::);
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    const uint32_t tmem = S.tmem_base;
    const int4* tiles = reinterpret_cast<const int4*>(a.tiles);

    auto local_of = [&](int slot) -> int {
        const int g = a.slot_client ? a.slot_client[slot] : a.first_client + slot;
        const int c = g - a.first_client;
        return (c >= 0 && c < a.n_local) ? c : -1;
    };
    constexpr uint32_t IDESC1 = idesc_tf32(128, 32), IDESC2 = idesc_tf32(128, 128);
    constexpr uint32_t XBYTES = (PREC == 3 ? 2u : 1u) * 4u * XBLK;
    uint32_t g = 0;               // tiles processed so far by this CTA: selects D1 buffer and barrier parities

    for (int u0 = blockIdx.x * unit; u0 < a.n_tiles; u0 += gridDim.x * unit) {
    const int t_end = min(a.n_tiles, u0 + unit);
    int t = u0;
    while (t < t_end) {
        // ---- a run = consecutive tiles of one slot (= one model) ----
        const int slot = tiles[t].x;
        int te = t + 1;
        for (;;) {                // warp-parallel scan for the end of the run (same result in every warp)
            const int q = te + lane;
            const bool same = q < t_end && tiles[q].x == slot;
            const unsigned m = __ballot_sync(0xffffffffu, same);
            if (m == 0xffffffffu) { te += 32; continue; }
            te += __ffs(~m) - 1;
            break;
        }
        const int c = local_of(slot);
        if (c < 0) { t = te; continue; }             // model lives on another rank: its owner scores it
        __syncthreads();                             // previous run fully drained (its epilogues waited on every MMA)
        stage_model<PREC>(S, a.slab + (size_t)c * P_PAD);

        if (warp == 4) {
            if (elect_one()) {      // elect.sync: ptxas emits the UTCHMMA sequence without per-MMA election loops
                auto load_tile = [&](int ti) {
                    const int row0 = tiles[ti].y;
                    mbar_expect_tx(&S.bar_x, XBYTES);
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb) {
                        tma_load_2d(S.x + kb * XBLK, &xmap, kb * 32, row0, &S.bar_x);
                        if (PREC == 3) tma_load_2d(S.xlo + kb * XBLK, &xlomap, kb * 32, row0, &S.bar_x);
                    }
                };
                load_tile(t);
                uint32_t gg = g;
                for (int ti = t; ti < te; ++ti, ++gg) {
                    const uint32_t b = gg & 1u, ph = gg & 1u, ph1 = (gg >> 1) & 1u;
                    mbar_wait(&S.bar_x, ph);
                    tc_fence_after();
                    const uint32_t d1 = tmem + (b ? COL_D1B : COL_D1A);
#pragma unroll
                    for (int kb = 0; kb < 4; ++kb) {
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t ad = smem_desc_sw128(S.x + kb * XBLK + ks * 32);
                            const uint64_t bd = smem_desc_sw128(S.w1 + kb * W1BLK + ks * 32);
                            mma_ss(d1, ad, bd, IDESC1, (kb | ks) ? 1u : 0u);
                            if (PREC == 3) {
                                mma_ss(d1, ad, smem_desc_sw128(S.w1lo + kb * W1BLK + ks * 32), IDESC1, 1u);
                                mma_ss(d1, smem_desc_sw128(S.xlo + kb * XBLK + ks * 32), bd, IDESC1, 1u);
                            }
                        }
                    }
                    tc_commit(&S.bar_d1[b]);
                    mbar_wait(&S.bar_d1[b], ph1);           // MMA 1 has consumed the x tile: refill it now
                    if (ti + 1 < te) load_tile(ti + 1);
                    mbar_wait(&S.bar_h2, ph);               // epilogue 1 done with D1 (and h2 is in TMEM)
                    if (need_sse) {
                        tc_fence_after();
#pragma unroll
                        for (int ks = 0; ks < 4; ++ks) {
                            const uint64_t bd = smem_desc_sw128(S.w4 + ks * 32);
                            mma_ts(tmem + COL_D2, tmem + COL_A2 + ks * 8, bd, IDESC2, ks ? 1u : 0u);
                            if (PREC == 3) {
                                mma_ts(tmem + COL_D2, tmem + COL_A2 + ks * 8, smem_desc_sw128(S.w4lo + ks * 32), IDESC2, 1u);
                                mma_ts(tmem + COL_D2, tmem + COL_A2LO + ks * 8, bd, IDESC2, 1u);
                            }
                        }
                        tc_commit(&S.bar_d2);
                    }
                }
            }
            __syncwarp();
        } else {
            const int r = warp * 32 + lane;                      // tile row = TMEM lane
            const uint32_t lane_base = tmem + ((uint32_t)(warp * 32) << 16);
            uint32_t gg = g;
            for (int ti = t; ti < te; ++ti, ++gg) {
                const int4 tl = tiles[ti];                        // {slot, row0, nrows, out0}
                const uint32_t b = gg & 1u, ph = gg & 1u, ph1 = (gg >> 1) & 1u;
                // ---- epilogue 1: one thread per row ----
                mbar_wait(&S.bar_d1[b], ph1);
                tc_fence_after();
                uint32_t v[32];
                tmem_ld32(lane_base + (b ? COL_D1B : COL_D1A), v);
                float h1[28];
#pragma unroll
                for (int j = 0; j < H; ++j) h1[j] = fmaxf(__uint_as_float(v[j]) + S.b1[j], 0.f);
                h1[27] = 0.f;
                float z[8];
#pragma unroll
                for (int i = 0; i < Z; ++i) {
                    float acc = S.b2[i];
#pragma unroll
                    for (int j4 = 0; j4 < 28; j4 += 4) {
                        const float4 w = *reinterpret_cast<const float4*>(&S.w2[i * 28 + j4]);
                        acc = fmaf(h1[j4], w.x, fmaf(h1[j4 + 1], w.y, fmaf(h1[j4 + 2], w.z, fmaf(h1[j4 + 3], w.w, acc))));
                    }
                    z[i] = acc;
                }
                z[7] = 0.f;
                if (a.out_lat && r < tl.z) {
                    float4* o = reinterpret_cast<float4*>(a.out_lat + (size_t)(tl.w + r) * 8);
                    o[0] = make_float4(z[0], z[1], z[2], z[3]);
                    o[1] = make_float4(z[4], z[5], z[6], 0.f);
                }
                if (!need_sse) {
                    tc_fence_before();
                    mbar_arrive(&S.bar_h2);                      // D1 buffer may be reused
                    continue;
                }
#pragma unroll
                for (int j = 0; j < H; ++j) {
                    const float4 wa = *reinterpret_cast<const float4*>(&S.w3[j * 8]);
                    const float4 wb = *reinterpret_cast<const float4*>(&S.w3[j * 8 + 4]);
                    float acc = S.b3[j];
                    acc = fmaf(z[0], wa.x, fmaf(z[1], wa.y, fmaf(z[2], wa.z, fmaf(z[3], wa.w, acc))));
                    acc = fmaf(z[4], wb.x, fmaf(z[5], wb.y, fmaf(z[6], wb.z, acc)));
                    const float hv = fmaxf(acc, 0.f);
                    if (PREC == 3) { v[j] = trunc_tf32(hv); h1[j] = hv - __uint_as_float(v[j]); }   // h1[] reused: lo part
                    else v[j] = cvt_tf32(hv);
                }
                v[27] = __float_as_uint(1.0f);                   // constant-one column: folds b4 into MMA 2
#pragma unroll
                for (int j = H + 1; j < 32; ++j) v[j] = 0u;
                tmem_st32(lane_base + COL_A2, v);
                if (PREC == 3) {
#pragma unroll
                    for (int j = 0; j < H; ++j) v[j] = __float_as_uint(h1[j]);
                    v[27] = 0u;
                    tmem_st32(lane_base + COL_A2LO, v);
                }
                tc_fence_before();
                mbar_arrive(&S.bar_h2);
                // ---- epilogue 2: squared reconstruction error; x comes straight from global/L2 ----
                const long long grow = (long long)tl.y + r;
                const bool live = r < tl.z && grow < x_rows;
                const float4* xg = reinterpret_cast<const float4*>(a.X + (size_t)(live ? grow : 0) * XP);
                float4 xa[8];
#pragma unroll
                for (int q = 0; q < 8; ++q) xa[q] = live ? __ldg(xg + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                mbar_wait(&S.bar_d2, ph);
                tc_fence_after();
                float sse = 0.f;
#pragma unroll
                for (int cb = 0; cb < 4; ++cb) {
                    float4 xn[8];
                    if (cb < 3) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) xn[q] = live ? __ldg(xg + (cb + 1) * 8 + q) : make_float4(0.f, 0.f, 0.f, 0.f);
                    }
                    tmem_ld32(lane_base + COL_D2 + cb * 32, v);
#pragma unroll
                    for (int q = 0; q < 8; ++q) {
                        const int d0 = cb * 32 + q * 4;
                        const float e0 = __uint_as_float(v[q * 4 + 0]) - xa[q].x, e1 = __uint_as_float(v[q * 4 + 1]) - xa[q].y;
                        const float e2 = __uint_as_float(v[q * 4 + 2]) - xa[q].z, e3 = __uint_as_float(v[q * 4 + 3]) - xa[q].w;
                        if (d0 + 0 < D) sse = fmaf(e0, e0, sse);
                        if (d0 + 1 < D) sse = fmaf(e1, e1, sse);
                        if (d0 + 2 < D) sse = fmaf(e2, e2, sse);
                        if (d0 + 3 < D) sse = fmaf(e3, e3, sse);
                    }
                    if (cb < 3) {
#pragma unroll
                        for (int q = 0; q < 8; ++q) xa[q] = xn[q];
                    }
                }
                if (r < tl.z) a.out_sse[tl.w + r] = sse;
            }
        }
        g += (uint32_t)(te - t);
        t = te;
    }
    }
    // teardown: everyone done with tensor memory, then the allocating warp frees it
    tc_fence_before();
    __syncthreads();
    if (warp == 4) asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;" ::"r"(tmem), "r"(TMEM_COLS));
}

}  // namespace tc
}  // namespace fm

// ---- host side: tensor map + launch ---------------------------------------------------------
namespace {
typedef CUresult (*EncodeTiledFn)(CUtensorMap*, CUtensorMapDataType, cuuint32_t, void*, const cuuint64_t*,
                                  const cuuint64_t*, const cuuint32_t*, const cuuint32_t*, CUtensorMapInterleave,
                                  CUtensorMapSwizzle, CUtensorMapL2promotion, CUtensorMapFloatOOBfill);
EncodeTiledFn get_encode() {
    static EncodeTiledFn fn = nullptr;
    if (!fn) {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q;
        if (cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q) == cudaSuccess &&
            q == cudaDriverEntryPointSuccess)
            fn = reinterpret_cast<EncodeTiledFn>(p);
    }
    return fn;
}
}  // namespace

static int encode_map(EncodeTiledFn enc, CUtensorMap* map, const float* base, int64_t rows) {
    const cuuint64_t gdim[2] = {(cuuint64_t)fm::XP, (cuuint64_t)rows};
    const cuuint64_t gstride[1] = {(cuuint64_t)fm::XP * sizeof(float)};
    const cuuint32_t box[2] = {32u, (cuuint32_t)fm::tc::TROWS};
    const cuuint32_t estride[2] = {1u, 1u};
    CUresult r = enc(map, CU_TENSOR_MAP_DATA_TYPE_FLOAT32, 2, const_cast<float*>(base), gdim, gstride, box, estride,
                     CU_TENSOR_MAP_INTERLEAVE_NONE, CU_TENSOR_MAP_SWIZZLE_128B, CU_TENSOR_MAP_L2_PROMOTION_L2_128B,
                     CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    return r == CUDA_SUCCESS ? 0 : (int)cudaErrorInvalidValue;
}

extern "C" int fm_split_residual(const float* x, float* x_lo, int64_t rows, void* stream) {
    if (rows <= 0) return 0;
    const size_t n4 = (size_t)rows * fm::XP / 4;
    const int grid = (int)((n4 + 255) / 256 < 1184 ? (n4 + 255) / 256 : 1184);
    fm::tc::split_residual_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(reinterpret_cast<const float4*>(x),
                                                                          reinterpret_cast<float4*>(x_lo), n4);
    return (int)cudaGetLastError();
}

// x_rows = number of valid rows behind args->X (rows past it are zero-filled by the TMA unit).
// precision 3 needs x_lo = the residual shard of X produced by fm_split_residual (same shape as X).
extern "C" int fm_score_rows_tc(const FmScoreArgs* args, int64_t x_rows, int32_t precision, const float* x_lo, void* stream) {
    using namespace fm;
    using namespace fm::tc;
    if (args->n_tiles <= 0) return 0;
    if (precision != 1 && precision != 3) return (int)cudaErrorInvalidValue;
    const bool x3 = precision == 3;
    if (x3 && !x_lo) return (int)cudaErrorInvalidValue;
    EncodeTiledFn enc = get_encode();
    if (!enc) return (int)cudaErrorNotSupported;
    CUtensorMap map, lomap;
    int rc = encode_map(enc, &map, args->X, x_rows);
    if (rc) return rc;
    rc = encode_map(enc, &lomap, x3 ? x_lo : args->X, x_rows);
    if (rc) return rc;
    static int sms = 0;
    if (!sms) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const size_t smem = (x3 ? sizeof(TcSmem<3>) : sizeof(TcSmem<1>)) + 1024;
    cudaError_t err = x3 ? cudaFuncSetAttribute(score_rows_tc_kernel<3>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem)
                         : cudaFuncSetAttribute(score_rows_tc_kernel<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    const int max_cta = (x3 ? 1 : 2) * sms;
    int per = (args->n_tiles + max_cta - 1) / max_cta;        // contiguous share ...
    if (per > 32) per = 32;                                     // ... capped: units of <= 32 tiles, dealt round-robin
    int grid = (args->n_tiles + per - 1) / per;
    if (grid > max_cta) grid = max_cta;
    if (x3) score_rows_tc_kernel<3><<<grid, NT_TC, smem, (cudaStream_t)stream>>>(*args, map, lomap, (long long)x_rows, per);
    else score_rows_tc_kernel<1><<<grid, NT_TC, smem, (cudaStream_t)stream>>>(*args, map, lomap, (long long)x_rows, per);
    return (int)cudaGetLastError();
}
