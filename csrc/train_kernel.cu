// T1 -- persistent per-client local training (fp32; the three batch GEMMs on warp-level tensor-core MMA, 3xTF32).
//
// One CTA owns one selected client for its whole ClientTrainer.run() call
// (reference: src/Trainer/client_trainer.py:360-419): weights (27 KB), Adam moments (54 KB) and,
// for FedProx, the anchor (27 KB) live in shared memory across all epochs and steps; the host is
// not involved between steps.  Batches of <= BT rows are streamed global->shared by the TMA unit
// (1-D bulk async copies completing on an mbarrier, double buffered) while the previous step
// computes.  Per step: 4-layer forward, SAE/AE loss, backward, FedProx gradient and Adam, all
// fused; weight gradients are consumed by the Adam update in registers and never stored.
//
// The batch is 12 rows and every step depends on the previous step's weights, so the
// matrix products are [12 x K] x [K x N] with N <= 115: far below a tcgen05 tile (M >= 64, and one
// thread per TMEM lane in the epilogue).  The m16n8k8 warp-level MMA fits exactly: the batch is the
// M = 16 tile.  Layer 1, the output layer (+ loss) and the data gradient run as error-compensated
// 3xTF32 mma.sync (hi*hi + hi*lo + lo*hi, fp32 accumulate: fp32-grade, same trick as the scoring
// kernels); operands come from shared memory with ldmatrix (pitches chosen conflict-free).  The
// bottleneck 27 -> 7 -> 27, the weight gradients and Adam stay on FFMA.  FM_TRAIN_MMA=0 in the
// environment selects the all-FFMA variant (also used for batches of 13..32 rows).
#include <cstdlib>
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

// phase timing (debug): thread 0 of CTA 0 accumulates SM cycles per phase when args.phase_cycles != NULL
#define FM_TICK(slot) do { if (dbg) { dbg_sink += *reinterpret_cast<volatile int*>(&S.flag[0]); /* forces the deferred barrier wait */ const long long now_ = clock64(); dbg_acc[slot] += now_ - dbg_t; dbg_t = now_; } } while (0)

// same clock, no forced barrier wait: splits a phase into "this warp's work" and "waiting for the others"
#define FM_TICK_WORK(slot) do { if (dbg) { const long long now_ = clock64(); dbg_acc[slot] += now_ - dbg_t; dbg_t = now_; } } while (0)

constexpr int NT = 512;  // threads per CTA

// shared-memory weight layout: every matrix row padded to a multiple of 4 floats so rows can be
// read with LDS.128, pitches chosen so 8 consecutive rows hit 8 distinct 16-B bank groups.
constexpr int W1P = 116, W2P = 28, W3P = 8, W4P = 28;
constexpr int S_W1 = 0;
constexpr int S_B1 = S_W1 + H * W1P;
constexpr int S_W2 = S_B1 + 28;
constexpr int S_B2 = S_W2 + Z * W2P;
constexpr int S_W3 = S_B2 + 8;
constexpr int S_B3 = S_W3 + H * W3P;
constexpr int S_W4 = S_B3 + 28;
constexpr int S_B4 = S_W4 + D * W4P;
constexpr int S_TOT = S_B4 + 116;  // 6944 floats
static_assert(S_TOT % 4 == 0, "");

__device__ __forceinline__ int sidx(int p) {
    if (p < OFF_B1) { int j = p / D; return S_W1 + j * W1P + (p - j * D); }
    if (p < OFF_W2) return S_B1 + (p - OFF_B1);
    if (p < OFF_B2) { int q = p - OFF_W2, i = q / H; return S_W2 + i * W2P + (q - i * H); }
    if (p < OFF_W3) return S_B2 + (p - OFF_B2);
    if (p < OFF_B3) { int q = p - OFF_W3, j = q / Z; return S_W3 + j * W3P + (q - j * Z); }
    if (p < OFF_W4) return S_B3 + (p - OFF_B3);
    if (p < OFF_B4) { int q = p - OFF_W4, d = q / H; return S_W4 + d * W4P + (q - d * H); }
    return S_B4 + (p - OFF_B4);
}

// C[b][n] = act(bias[n] + sum_k A[b][k] W[n][k]); K = 4*K4 (zero padded), rows b < BT.
// item = (n, group of 4 rows); KS adjacent lanes split the K chunks and shuffle-reduce.
template <int BT, int N, int K4, int KS, bool RELU>
__device__ __forceinline__ void layer_nt(const float* __restrict__ A, int AP, const float* __restrict__ W,
                                         int WP, const float* __restrict__ bias, float* __restrict__ C,
                                         int CP) {
    constexpr int ITEMS = N * (BT / 4);
    constexpr int PER_PASS = NT / KS;
    const int s = threadIdx.x % KS;
    for (int base = 0; base < ITEMS; base += PER_PASS) {
        const int it = base + threadIdx.x / KS;
        const bool active = it < ITEMS;
        const int n = active ? it % N : 0, bg = active ? it / N : 0;
        float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
        if (active) {
            const float* wrow = W + n * WP;
            const float* arow = A + (bg * 4) * AP;
#pragma unroll 2
            for (int c = s; c < K4; c += KS) {
                const float4 w = *reinterpret_cast<const float4*>(wrow + 4 * c);
                const float4 a0 = *reinterpret_cast<const float4*>(arow + 4 * c);
                const float4 a1 = *reinterpret_cast<const float4*>(arow + AP + 4 * c);
                const float4 a2 = *reinterpret_cast<const float4*>(arow + 2 * AP + 4 * c);
                const float4 a3 = *reinterpret_cast<const float4*>(arow + 3 * AP + 4 * c);
                acc0 = fmaf(a0.x, w.x, fmaf(a0.y, w.y, fmaf(a0.z, w.z, fmaf(a0.w, w.w, acc0))));
                acc1 = fmaf(a1.x, w.x, fmaf(a1.y, w.y, fmaf(a1.z, w.z, fmaf(a1.w, w.w, acc1))));
                acc2 = fmaf(a2.x, w.x, fmaf(a2.y, w.y, fmaf(a2.z, w.z, fmaf(a2.w, w.w, acc2))));
                acc3 = fmaf(a3.x, w.x, fmaf(a3.y, w.y, fmaf(a3.z, w.z, fmaf(a3.w, w.w, acc3))));
            }
        }
#pragma unroll
        for (int o = 1; o < KS; o <<= 1) {
            acc0 += __shfl_xor_sync(0xffffffffu, acc0, o);
            acc1 += __shfl_xor_sync(0xffffffffu, acc1, o);
            acc2 += __shfl_xor_sync(0xffffffffu, acc2, o);
            acc3 += __shfl_xor_sync(0xffffffffu, acc3, o);
        }
        if (active && s == 0) {
            const float bb = bias[n];
            float r0 = acc0 + bb, r1 = acc1 + bb, r2 = acc2 + bb, r3 = acc3 + bb;
            if (RELU) { r0 = fmaxf(r0, 0.f); r1 = fmaxf(r1, 0.f); r2 = fmaxf(r2, 0.f); r3 = fmaxf(r3, 0.f); }
            float* c = C + (bg * 4) * CP + n;
            c[0] = r0; c[CP] = r1; c[2 * CP] = r2; c[3 * CP] = r3;
        }
    }
}

// Data gradient: GinT[k][b] = mask * sum_n GoutT[n][b] * W[n][k]; gradients are stored transposed
// ([feature][BT]) so four batch rows are one LDS.128.  mask = (Act[b][k] > 0) when RELU_MASK.
template <int BT, int N, int K, int KS, bool RELU_MASK>
__device__ __forceinline__ void layer_dgrad(const float* __restrict__ GoutT, const float* __restrict__ W, int WP,
                                            const float* __restrict__ Act, int ActP, float* __restrict__ GinT) {
    constexpr int ITEMS = K * (BT / 4);
    constexpr int PER_PASS = NT / KS;
    const int s = threadIdx.x % KS;
    for (int base = 0; base < ITEMS; base += PER_PASS) {
        const int it = base + threadIdx.x / KS;
        const bool active = it < ITEMS;
        const int k = active ? it % K : 0, bg = active ? it / K : 0;
        float4 acc = make_float4(0.f, 0.f, 0.f, 0.f);
        if (active) {
#pragma unroll 4
            for (int n = s; n < N; n += KS) {
                const float w = W[n * WP + k];
                const float4 g = *reinterpret_cast<const float4*>(GoutT + n * BT + bg * 4);
                acc.x = fmaf(g.x, w, acc.x); acc.y = fmaf(g.y, w, acc.y);
                acc.z = fmaf(g.z, w, acc.z); acc.w = fmaf(g.w, w, acc.w);
            }
        }
#pragma unroll
        for (int o = 1; o < KS; o <<= 1) {
            acc.x += __shfl_xor_sync(0xffffffffu, acc.x, o);
            acc.y += __shfl_xor_sync(0xffffffffu, acc.y, o);
            acc.z += __shfl_xor_sync(0xffffffffu, acc.z, o);
            acc.w += __shfl_xor_sync(0xffffffffu, acc.w, o);
        }
        if (active && s == 0) {
            if (RELU_MASK) {
                const float* a = Act + (bg * 4) * ActP + k;
                acc.x = a[0] > 0.f ? acc.x : 0.f;
                acc.y = a[ActP] > 0.f ? acc.y : 0.f;
                acc.z = a[2 * ActP] > 0.f ? acc.z : 0.f;
                acc.w = a[3 * ActP] > 0.f ? acc.w : 0.f;
            }
            *reinterpret_cast<float4*>(GinT + k * BT + bg * 4) = acc;
        }
    }
}

// ---- warp-level tensor-core helpers (m16n8k8, TF32 inputs, fp32 accumulate) ---------------------
// fragment coordinates, g = lane >> 2, t = lane & 3:  A: a0 (g, t) a1 (g+8, t) a2 (g, t+4) a3 (g+8, t+4);
// B (k, n): b0 (t, g) b1 (t+4, g);  C: c0 (g, 2t) c1 (g, 2t+1) c2 (g+8, 2t) c3 (g+8, 2t+1).
__device__ __forceinline__ void ldsm4(uint32_t (&r)[4], const float* row_ptr) {   // four 8 x (4 floats) blocks
    asm volatile("ldmatrix.sync.aligned.m8n8.x4.shared.b16 {%0, %1, %2, %3}, [%4];"
                 : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]) : "r"(smem_u32(row_ptr)));
}
__device__ __forceinline__ void mma_tf32(float (&c)[4], const uint32_t (&a)[4], uint32_t b0, uint32_t b1) {
    asm volatile("mma.sync.aligned.m16n8k8.row.col.f32.tf32.tf32.f32 {%0, %1, %2, %3}, {%4, %5, %6, %7}, {%8, %9}, {%0, %1, %2, %3};"
                 : "+f"(c[0]), "+f"(c[1]), "+f"(c[2]), "+f"(c[3])
                 : "r"(a[0]), "r"(a[1]), "r"(a[2]), "r"(a[3]), "r"(b0), "r"(b1));
}
// C += A B in 3xTF32: A = ah + al, B = bh + bl with ah/bh the fp32 values themselves (the tensor core ignores their
// low 13 mantissa bits, i.e. sees trunc_tf32) and al/bl the exact fp32 remainders, which are kept in shared memory
// next to their operands (tf32_lo below) -- the MMA loops contain no arithmetic, only ldmatrix and mma.sync.
// Work split (measured with scripts/mma_sync_probe.cu: an mma.sync issues every 8 cycles per SM sub-partition with a
// 20-cycle dependent latency; an ldmatrix.x4 costs 4 cycles of the SM's shared-memory bandwidth): eight warps, each
// owning two adjacent 8-column output tiles (they share the A fragments, one ldmatrix.x4 fetches the B fragments of
// both) and, for the two long-K products, one quarter of the k-steps.  The partial sums of the k-quarters meet in
// shared memory and are added by the consumer of the result (the warp-per-row bottleneck code), so no extra block
// barrier is needed.  Per phase: 128 ldmatrix (512 cycles of smem bandwidth) and 48 MMAs per sub-partition (420).
struct Acc2 {
    float hh[4], sm[4];             // head product | the two small cross terms
    __device__ __forceinline__ void zero() {
#pragma unroll
        for (int i = 0; i < 4; ++i) { hh[i] = 0.f; sm[i] = 0.f; }
    }
    __device__ __forceinline__ float sum(int i) const { return sm[i] + hh[i]; }
};
// one k-step for two adjacent 8-column output tiles that share the A fragments; b* = {tile0 b0, tile0 b1, tile1 b0, tile1 b1}.
// Issue order keeps four MMAs between two writes of the same accumulator.
__device__ __forceinline__ void mma3x2(Acc2& c0, Acc2& c1, const uint32_t (&ah)[4], const uint32_t (&al)[4],
                                       const uint32_t (&bh)[4], const uint32_t (&bl)[4]) {
    mma_tf32(c0.hh, ah, bh[0], bh[1]);
    mma_tf32(c1.hh, ah, bh[2], bh[3]);
    mma_tf32(c0.sm, ah, bl[0], bl[1]);
    mma_tf32(c1.sm, ah, bl[2], bl[3]);
    mma_tf32(c0.sm, al, bh[0], bh[1]);
    mma_tf32(c1.sm, al, bh[2], bh[3]);
}
__device__ __forceinline__ float tf32_lo(float x) { return x - __uint_as_float(__float_as_uint(x) & 0xFFFFE000u); }

struct AdamStep {
    float step_size;     // lr / (1 - b1^t)
    float inv_sqrt_bc2;  // 1 / sqrt(1 - b2^t)
    float mu;            // FedProx
};

__device__ __forceinline__ void adam1(float& w, float& m, float& v, float anchor, float g, const AdamStep& st,
                                      float& prox_acc) {
    if (st.mu != 0.f) {
        const float dl = w - anchor;
        prox_acc = fmaf(dl, dl, prox_acc);
        g = fmaf(2.f * st.mu, dl, g);
    }
    m = fmaf(g - m, 0.1f, m);                     // exp_avg.lerp_(g, 1 - beta1)
    v = fmaf(0.001f * g, g, v * 0.999f);          // exp_avg_sq.mul_(beta2).addcmul_(g, g, 1 - beta2)
    float sq;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq) : "f"(v));
    const float denom = fmaf(sq, st.inv_sqrt_bc2, 1e-8f);
    w = fmaf(-st.step_size, __fdividef(m, denom), w);
}

// ---- packed fp32 pairs (FFMA2 / FMUL2 / FADD2 on sm_100): two IEEE operations per issue slot, same rounding ------
typedef unsigned long long f32x2;
__device__ __forceinline__ f32x2 pk2(float a, float b) { f32x2 r; asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b)); return r; }
__device__ __forceinline__ f32x2 bc2(float a) { return pk2(a, a); }
__device__ __forceinline__ float2 up2(f32x2 r) { float2 d; asm("mov.b64 {%0, %1}, %2;" : "=f"(d.x), "=f"(d.y) : "l"(r)); return d; }
__device__ __forceinline__ f32x2 fma2(f32x2 a, f32x2 b, f32x2 c) { f32x2 d; asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c)); return d; }
__device__ __forceinline__ f32x2 mul2(f32x2 a, f32x2 b) { f32x2 d; asm("mul.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }
__device__ __forceinline__ f32x2 sub2(f32x2 a, f32x2 b) { f32x2 d; asm("sub.rn.f32x2 %0, %1, %2;" : "=l"(d) : "l"(a), "l"(b)); return d; }

// adam1 on two parameters at once: the same operations in the same order, packed
__device__ __forceinline__ void adam2(float& w0, float& w1, float& m0, float& m1, float& v0, float& v1, float an0, float an1,
                                      f32x2 g, const AdamStep& st, float& prox_acc) {
    f32x2 w = pk2(w0, w1), m = pk2(m0, m1), v = pk2(v0, v1);
    if (st.mu != 0.f) {
        const f32x2 dl = sub2(w, pk2(an0, an1));
        const float2 d = up2(dl);
        prox_acc = fmaf(d.x, d.x, prox_acc);
        prox_acc = fmaf(d.y, d.y, prox_acc);
        g = fma2(bc2(2.f * st.mu), dl, g);
    }
    m = fma2(sub2(g, m), bc2(0.1f), m);
    v = fma2(mul2(bc2(0.001f), g), g, mul2(v, bc2(0.999f)));
    const float2 vv = up2(v);
    float sq0, sq1;
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq0) : "f"(vv.x));
    asm("sqrt.approx.ftz.f32 %0, %1;" : "=f"(sq1) : "f"(vv.y));
    const float2 den = up2(fma2(pk2(sq0, sq1), bc2(st.inv_sqrt_bc2), bc2(1e-8f)));
    const float2 mm = up2(m);
    w = fma2(bc2(-st.step_size), pk2(__fdividef(mm.x, den.x), __fdividef(mm.y, den.y)), w);
    const float2 ww = up2(w);
    w0 = ww.x; w1 = ww.y; m0 = mm.x; m1 = mm.y; v0 = vv.x; v1 = vv.y;
}

__device__ __forceinline__ void adam_update(float* __restrict__ Ws, float* __restrict__ Ms, float* __restrict__ Vs,
                                            const float* __restrict__ As, int idx, float g, const AdamStep& st,
                                            float& prox_acc) {
    float w = Ws[idx], m = Ms[idx], v = Vs[idx];
    adam1(w, m, v, st.mu != 0.f ? As[idx] : 0.f, g, st, prox_acc);
    Ws[idx] = w; Ms[idx] = m; Vs[idx] = v;
}

// four consecutive parameters at a 16-B aligned index: vector LDS/STS, conflict-free across lanes.
// Padding slots need no mask: their gradient is exactly 0 (zero-padded activations), so w, m, v stay 0.
__device__ __forceinline__ void adam_update4(float* __restrict__ Ws, float* __restrict__ Ms, float* __restrict__ Vs,
                                             const float* __restrict__ As, int idx, float4 g, const AdamStep& st,
                                             float& prox_acc) {
    float4 w = *reinterpret_cast<float4*>(Ws + idx);
    float4 m = *reinterpret_cast<float4*>(Ms + idx);
    float4 v = *reinterpret_cast<float4*>(Vs + idx);
    float4 an = make_float4(0.f, 0.f, 0.f, 0.f);
    if (st.mu != 0.f) an = *reinterpret_cast<const float4*>(As + idx);
    adam1(w.x, m.x, v.x, an.x, g.x, st, prox_acc);
    adam1(w.y, m.y, v.y, an.y, g.y, st, prox_acc);
    adam1(w.z, m.z, v.z, an.z, g.z, st, prox_acc);
    adam1(w.w, m.w, v.w, an.w, g.w, st, prox_acc);
    *reinterpret_cast<float4*>(Ws + idx) = w;
    *reinterpret_cast<float4*>(Ms + idx) = m;
    *reinterpret_cast<float4*>(Vs + idx) = v;
}

// Weight gradient + Adam for one linear layer: dW[n][k] = sum_b GT[n][b] * A[b][k], dbias[n] = sum_b GT[n][b].
// item = (n, chunk of 4 consecutive k); the gradient never leaves registers.
template <int BT, int N, int K, int K4, int STRIDE>
__device__ __forceinline__ void wgrad_adam(int item0, const float* __restrict__ GT, const float* __restrict__ A,
                                           int AP, int sW, int WP, int sB, float* Ws, float* Ms, float* Vs,
                                           const float* As, const AdamStep& st, float& prox_acc) {
    constexpr int ITEMS = N * K4 + (N + 3) / 4;
    for (int it = item0; it < ITEMS; it += STRIDE) {
        if (it < N * K4) {
            const int n = it / K4, kc = it - n * K4;
            float4 g = make_float4(0.f, 0.f, 0.f, 0.f);
            const float* gt = GT + n * BT;
            const float* a = A + 4 * kc;
#pragma unroll
            for (int b4 = 0; b4 < BT; b4 += 4) {
                const float4 gg = *reinterpret_cast<const float4*>(gt + b4);
                const float4 a0 = *reinterpret_cast<const float4*>(a + (b4 + 0) * AP);
                const float4 a1 = *reinterpret_cast<const float4*>(a + (b4 + 1) * AP);
                const float4 a2 = *reinterpret_cast<const float4*>(a + (b4 + 2) * AP);
                const float4 a3 = *reinterpret_cast<const float4*>(a + (b4 + 3) * AP);
                g.x = fmaf(gg.x, a0.x, fmaf(gg.y, a1.x, fmaf(gg.z, a2.x, fmaf(gg.w, a3.x, g.x))));
                g.y = fmaf(gg.x, a0.y, fmaf(gg.y, a1.y, fmaf(gg.z, a2.y, fmaf(gg.w, a3.y, g.y))));
                g.z = fmaf(gg.x, a0.z, fmaf(gg.y, a1.z, fmaf(gg.z, a2.z, fmaf(gg.w, a3.z, g.z))));
                g.w = fmaf(gg.x, a0.w, fmaf(gg.y, a1.w, fmaf(gg.z, a2.w, fmaf(gg.w, a3.w, g.w))));
            }
            adam_update4(Ws, Ms, Vs, As, sW + n * WP + 4 * kc, g, st, prox_acc);
        } else {
            const int n0 = (it - N * K4) * 4;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const int n = n0 + q;
                if (n < N) {
                    float g = 0.f;
#pragma unroll
                    for (int b = 0; b < BT; ++b) g += GT[n * BT + b];
                    adam_update(Ws, Ms, Vs, As, sB + n, g, st, prox_acc);
                }
            }
        }
    }
}


// ---- the "small" role of the update phase: W2, W3 and every bias on 64 threads -------------------------------
// Measured: these two warps (one per SM sub-partition, nobody to hide latency behind) were the critical path of the
// phase -- 3.0 k cycles on their own -- because the three per-layer loops ran strictly one after the other: every
// Adam store to the weight array orders the next section's loads behind it.  Here all gradients are formed first
// (reads only, independent chains), then all weights / moments are loaded, updated and stored.
template <int BT>
__device__ __forceinline__ float4 wgrad_quad(const float* __restrict__ gt, const float* __restrict__ a, int AP) {
    float4 g = make_float4(0.f, 0.f, 0.f, 0.f);        // g[c] = sum_b gt[b] * a[b][c]
#pragma unroll
    for (int b4 = 0; b4 < BT; b4 += 4) {
        const float4 gg = *reinterpret_cast<const float4*>(gt + b4);
        const float4 a0 = *reinterpret_cast<const float4*>(a + (b4 + 0) * AP);
        const float4 a1 = *reinterpret_cast<const float4*>(a + (b4 + 1) * AP);
        const float4 a2 = *reinterpret_cast<const float4*>(a + (b4 + 2) * AP);
        const float4 a3 = *reinterpret_cast<const float4*>(a + (b4 + 3) * AP);
        g.x = fmaf(gg.x, a0.x, fmaf(gg.y, a1.x, fmaf(gg.z, a2.x, fmaf(gg.w, a3.x, g.x))));
        g.y = fmaf(gg.x, a0.y, fmaf(gg.y, a1.y, fmaf(gg.z, a2.y, fmaf(gg.w, a3.y, g.y))));
        g.z = fmaf(gg.x, a0.z, fmaf(gg.y, a1.z, fmaf(gg.z, a2.z, fmaf(gg.w, a3.z, g.z))));
        g.w = fmaf(gg.x, a0.w, fmaf(gg.y, a1.w, fmaf(gg.z, a2.w, fmaf(gg.w, a3.w, g.w))));
    }
    return g;
}
template <int BT>
__device__ __forceinline__ float row_sum(const float* __restrict__ gt) {
    float acc = 0.f;
#pragma unroll
    for (int b4 = 0; b4 < BT; b4 += 4) {
        const float4 gg = *reinterpret_cast<const float4*>(gt + b4);
        acc += (gg.x + gg.y) + (gg.z + gg.w);
    }
    return acc;
}
template <int BT, int N>
__device__ __forceinline__ float4 bias_quad(const float* __restrict__ GT, int n0) {
    return make_float4(n0 + 0 < N ? row_sum<BT>(GT + (n0 + 0) * BT) : 0.f, n0 + 1 < N ? row_sum<BT>(GT + (n0 + 1) * BT) : 0.f,
                       n0 + 2 < N ? row_sum<BT>(GT + (n0 + 2) * BT) : 0.f, n0 + 3 < N ? row_sum<BT>(GT + (n0 + 3) * BT) : 0.f);
}

// Weight gradient + Adam for a 4(n) x 4(k) register tile of one big matrix (W1 or W4).  The thread owns
// these 16 parameters for the whole kernel: their Adam moments stay in registers (m, v), the activations
// are read once per tile instead of once per output feature, and the weights are touched once per step.
template <int BT, int N, int K4>
__device__ __forceinline__ void wgrad_adam_tile(int item, const float* __restrict__ GT, const float* __restrict__ A,
                                                int AP, int sW, int WP, float* Ws, float* Wlo, const float* As, const AdamStep& st,
                                                float (&m)[16], float (&v)[16], float& prox_acc) {
    const int nt = item / K4, kc = item - nt * K4;
    const int n0 = nt * 4, nvalid = min(4, N - n0);
    f32x2 g[8];                                   // g[q * 2 + h]: row q of the tile, columns 2h, 2h + 1
#pragma unroll
    for (int i = 0; i < 8; ++i) g[i] = pk2(0.f, 0.f);
    const float* a = A + 4 * kc;
#pragma unroll
    for (int b4 = 0; b4 < BT; b4 += 4) {
        const float4 a0 = *reinterpret_cast<const float4*>(a + (b4 + 0) * AP);
        const float4 a1 = *reinterpret_cast<const float4*>(a + (b4 + 1) * AP);
        const float4 a2 = *reinterpret_cast<const float4*>(a + (b4 + 2) * AP);
        const float4 a3 = *reinterpret_cast<const float4*>(a + (b4 + 3) * AP);
        const f32x2 a0l = pk2(a0.x, a0.y), a0h = pk2(a0.z, a0.w), a1l = pk2(a1.x, a1.y), a1h = pk2(a1.z, a1.w);
        const f32x2 a2l = pk2(a2.x, a2.y), a2h = pk2(a2.z, a2.w), a3l = pk2(a3.x, a3.y), a3h = pk2(a3.z, a3.w);
#pragma unroll
        for (int q = 0; q < 4; ++q) {
            if (q < nvalid) {
                const float4 gg = *reinterpret_cast<const float4*>(GT + (n0 + q) * BT + b4);
                const f32x2 gx = bc2(gg.x), gy = bc2(gg.y), gz = bc2(gg.z), gw = bc2(gg.w);
                g[q * 2 + 0] = fma2(gx, a0l, fma2(gy, a1l, fma2(gz, a2l, fma2(gw, a3l, g[q * 2 + 0]))));
                g[q * 2 + 1] = fma2(gx, a0h, fma2(gy, a1h, fma2(gz, a2h, fma2(gw, a3h, g[q * 2 + 1]))));
            }
        }
    }
#pragma unroll
    for (int q = 0; q < 4; ++q) {
        if (q < nvalid) {
            const int idx = sW + (n0 + q) * WP + 4 * kc;
            float4 w = *reinterpret_cast<float4*>(Ws + idx);
            float4 an = make_float4(0.f, 0.f, 0.f, 0.f);
            if (st.mu != 0.f) an = *reinterpret_cast<const float4*>(As + idx);
            adam2(w.x, w.y, m[q * 4 + 0], m[q * 4 + 1], v[q * 4 + 0], v[q * 4 + 1], an.x, an.y, g[q * 2 + 0], st, prox_acc);
            adam2(w.z, w.w, m[q * 4 + 2], m[q * 4 + 3], v[q * 4 + 2], v[q * 4 + 3], an.z, an.w, g[q * 2 + 1], st, prox_acc);
            *reinterpret_cast<float4*>(Ws + idx) = w;
            if (Wlo) *reinterpret_cast<float4*>(Wlo + idx) = make_float4(tf32_lo(w.x), tf32_lo(w.y), tf32_lo(w.z), tf32_lo(w.w));
        }
    }
}

// moments of a register tile <-> packed global rows (p = off + n * K + k)
template <int N, int K, int K4>
__device__ __forceinline__ void tile_moments_io(int item, int off, float* gm, float* gv, float (&m)[16], float (&v)[16],
                                                bool store) {
    const int nt = item / K4, kc = item - nt * K4;
#pragma unroll
    for (int q = 0; q < 4; ++q) {
#pragma unroll
        for (int cix = 0; cix < 4; ++cix) {
            const int n = nt * 4 + q, k = kc * 4 + cix;
            const bool ok = n < N && k < K;
            if (store) { if (ok) { gm[off + n * K + k] = m[q * 4 + cix]; gv[off + n * K + k] = v[q * 4 + cix]; } }
            else { m[q * 4 + cix] = ok ? gm[off + n * K + k] : 0.f; v[q * 4 + cix] = ok ? gv[off + n * K + k] : 0.f; }
        }
    }
}

constexpr int W1_TILES = ((H + 3) / 4) * 29;    // 7 x 29 = 203 register tiles
constexpr int W4_TILES = ((D + 3) / 4) * 7;     // 29 x 7 = 203
constexpr int ROLE_W1 = 0, ROLE_W4 = 224, ROLE_SMALL = 448;   // first thread of each role (warp aligned)
static_assert(W1_TILES <= ROLE_W4 - ROLE_W1 && W4_TILES <= ROLE_SMALL - ROLE_W4, "role ranges too small");

template <int BT, bool MMA>
struct TrainSmem {
    // MMA variant: 16-row operand tiles (rows BT..15 stay zero), pitches 132 / 36 floats so that the eight
    // 16-byte rows of an ldmatrix block land in eight different bank groups
    static constexpr int RB = MMA ? 16 : BT;
    static constexpr int XS = MMA ? 132 : XP;
    static constexpr int HP = MMA ? 36 : 28, ZP = 8;
    alignas(16) float W[S_TOT];
    alignas(16) float M[S_TOT];
    alignas(16) float V[S_TOT];
    alignas(16) float x[2][RB * XS];
    alignas(16) float h1[RB * HP];
    alignas(16) float z[BT * ZP];
    alignas(16) float h2[RB * HP];
    alignas(16) float dxr[MMA ? 16 * 132 : 4];   // d(loss)/d(xhat) row-major [b][d]: A operand of the data gradient
    // exact TF32 remainders (x - trunc_tf32(x)) of the tensor-core operands, same indexing as their heads
    alignas(16) float Wlo[MMA ? S_TOT + 32 : 4];  // W1 / W4 regions only (+32: operand tiles read a little past the end)
    alignas(16) float xlo[2][MMA ? 16 * 132 : 4];   // per batch buffer; filled one step ahead by warps that idle in the output layer
    alignas(16) float h2lo[MMA ? 16 * 36 : 4];
    alignas(16) float dxrlo[MMA ? 16 * 132 : 4];
    alignas(16) float part[MMA ? 4 * 16 * 32 : 4];   // partial sums of the four k-quarters: [kq][row][column]
    alignas(16) float dxT[D * BT];     // d(loss)/d(xhat), transposed [d][b]
    alignas(16) float dp2T[H * BT];    // grad wrt pre-activation of decoder hidden
    alignas(16) float dzT[Z * BT];
    alignas(16) float dp1T[H * BT];
    float znorm[BT];
    double red[34];
    float sctab[64];                    // Adam bias corrections (step size, 1/sqrt(1 - b2^t)) of the next 32 train steps
    int flag[4];
    alignas(8) uint64_t bar[2];
    // FedProx anchor (S_TOT floats) follows only when mu != 0
};

template <int BT, bool MMA>
__global__ void __launch_bounds__(NT, 1) train_clients_kernel(const FmTrainArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    using SM = TrainSmem<BT, MMA>;
    constexpr int XS = SM::XS;
    SM& S = *reinterpret_cast<SM*>(smem_raw);
    float* As = reinterpret_cast<float*>(smem_raw + ((sizeof(SM) + 15) / 16) * 16);  // valid iff mu != 0
    const int tid = threadIdx.x, lane = tid & 31, warp = tid >> 5;
    const int c = a.sel_local[blockIdx.x];
    const int g = a.first_client + c;
    const int tr0 = a.train_off[c], n_tr = a.train_off[c + 1] - tr0;
    const int va0 = a.valid_off[g], n_va = a.valid_off[g + 1] - va0;
    const int B = a.batch;
    const int nbt = (n_tr + B - 1) / B, nbv = (n_va + B - 1) / B;
    const int steps_per_epoch = nbt + nbv;
    const bool prox = a.fedprox_mu != 0.f;
    float* gp = a.params + (size_t)c * P_PAD;
    float* gm = a.adam_m + (size_t)c * P_PAD;
    float* gv = a.adam_v + (size_t)c * P_PAD;

    // ---- load state into shared memory -----------------------------------------------------
    for (int i = tid; i < S_TOT; i += NT) { S.W[i] = 0.f; S.M[i] = 0.f; S.V[i] = 0.f; if (prox) As[i] = 0.f; }
    for (int i = tid; i < 2 * SM::RB * XS; i += NT) (&S.x[0][0])[i] = 0.f;
    for (int i = tid; i < SM::RB * SM::HP; i += NT) { S.h1[i] = 0.f; S.h2[i] = 0.f; }
    if (MMA) {
        for (int i = tid; i < 16 * 132; i += NT) { S.dxr[i] = 0.f; S.dxrlo[i] = 0.f; S.xlo[0][i] = 0.f; S.xlo[1][i] = 0.f; }
        for (int i = tid; i < 16 * 36; i += NT) S.h2lo[i] = 0.f;
        for (int i = tid; i < S_TOT + 32; i += NT) S.Wlo[i] = 0.f;
    }
    for (int i = tid; i < BT * SM::ZP; i += NT) S.z[i] = 0.f;
    for (int i = tid; i < D * BT; i += NT) S.dxT[i] = 0.f;
    for (int i = tid; i < H * BT; i += NT) { S.dp2T[i] = 0.f; S.dp1T[i] = 0.f; }
    for (int i = tid; i < Z * BT; i += NT) S.dzT[i] = 0.f;
    if (tid == 0) { mbar_init(&S.bar[0], 1); mbar_init(&S.bar[1], 1); fence_mbar_init(); }
    __syncthreads();
    for (int p = tid; p < NP; p += NT) {
        const int s = sidx(p);
        S.W[s] = gp[p]; S.M[s] = gm[p]; S.V[s] = gv[p];
        if (prox) As[s] = a.anchor[(size_t)c * P_PAD + p];
    }
    if (MMA) {
        __syncthreads();
        for (int i = tid; i < S_TOT; i += NT) S.Wlo[i] = tf32_lo(S.W[i]);
    }
    float tm[16], tv[16];          // Adam moments of this thread's register tile (roles W1 / W4)
#pragma unroll
    for (int i = 0; i < 16; ++i) { tm[i] = 0.f; tv[i] = 0.f; }
    if (tid < ROLE_W4) { if (tid < W1_TILES) tile_moments_io<H, D, 29>(tid, OFF_W1, gm, gv, tm, tv, false); }
    else if (tid < ROLE_SMALL) { if (tid - ROLE_W4 < W4_TILES) tile_moments_io<D, H, 7>(tid - ROLE_W4, OFF_W4, gm, gv, tm, tv, false); }
    fence_proxy_async();
    __syncthreads();

    // batch stream: per epoch nbt train batches then nbv validation batches
    constexpr int ISSUE_WARP = 14;                // a warp without tensor-core work in layer 1
    auto issue = [&](int q_in_epoch, int buf) {   // called by every lane of warp ISSUE_WARP
        const float* src; int rows;
        if (q_in_epoch < nbt) { const int r0 = q_in_epoch * B; rows = min(B, n_tr - r0); src = a.x_train + (size_t)(tr0 + r0) * XP; }
        else { const int r0 = (q_in_epoch - nbt) * B; rows = min(B, n_va - r0); src = a.x_valid + (size_t)(va0 + r0) * XP; }
        const uint32_t bytes = (uint32_t)rows * XP * 4u;
        if (lane == 0) mbar_expect_tx(&S.bar[buf], bytes);
        __syncwarp();
        if (MMA) {       // padded shared-memory pitch: one 512-byte copy per row, one lane each
            if (lane < rows) bulk_g2s(&S.x[buf][lane * XS], src + (size_t)lane * XP, XP * 4u, &S.bar[buf]);
        } else if (lane == 0) bulk_g2s(&S.x[buf][0], src, bytes, &S.bar[buf]);
    };

    // Adam bias corrections are computed in double like the reference (torch computes them on the host), but FP64
    // divide + sqrt cost ~1k cycles on one thread -- longer than the tensor-core layer 1 they used to hide behind.
    // Warp 15 therefore tabulates them for 32 steps at a time, one step per lane.
    const int t_step0 = a.adam_t[c];
    int ts = 0;                                         // train steps taken in this launch (uniform)
    double b1t = 1.0, b2t = 1.0, p1 = 1.0, p2 = 1.0;    // warp 15: beta^t at the last tabulated step; beta^(lane+1)
    if (warp == 15) {
        b1t = pow(0.9, (double)t_step0); b2t = pow(0.999, (double)t_step0);
        p1 = pow(0.9, (double)(lane + 1)); p2 = pow(0.999, (double)(lane + 1));
    }
    float min_valid = __int_as_float(0x7f800000);
    int worse = 0, epochs_done = 0;
    uint32_t uses0 = 0, uses1 = 0;   // completed waits per buffer -> parity
    long long q = 0;                  // global step index (selects the buffer)
    bool outstanding = false;
    if (warp == ISSUE_WARP && steps_per_epoch > 0 && a.epochs > 0 && n_tr > 0) issue(0, 0);
    outstanding = (steps_per_epoch > 0 && a.epochs > 0 && n_tr > 0);
    const int n_epochs = (n_tr > 0) ? a.epochs : 0;

    const bool dbg = a.phase_cycles != nullptr && blockIdx.x == 0 && tid == 0;
    long long dbg_acc[16] = {0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0, 0};
    long long dbg_t = clock64();
    int dbg_sink = 0;
    for (int e = 0; e < n_epochs; ++e) {
        float loss_acc = 0.f;   // this thread's share of sum over train batches of batch loss
        float vloss_acc = 0.f;
        for (int qe = 0; qe < steps_per_epoch; ++qe, ++q) {
            const int buf = (int)(q & 1);
            const bool is_train = qe < nbt;
            const int r0 = is_train ? qe * B : (qe - nbt) * B;
            const int rows = min(B, (is_train ? n_tr : n_va) - r0);
            // S0: wait for this batch, prefetch the next one into the other buffer
            FM_TICK(7);
            mbar_wait(&S.bar[buf], (buf ? uses1 : uses0) & 1u);
            if (buf) ++uses1; else ++uses0;
            const bool has_next = (qe + 1 < steps_per_epoch) || (e + 1 < n_epochs);
            // (tensor-core variant: twelve row copies, ~800 cycles of issue on warp 14 -- hidden behind the layer-1 MMAs of
            // warps 0-7, there is no block barrier before those)
            if (warp == ISSUE_WARP && has_next) issue((qe + 1) % steps_per_epoch, buf ^ 1);
            outstanding = has_next;
            const float* X = &S.x[buf][0];
            FM_TICK(0);
            if (is_train && warp == 15 && (ts & 31) == 0) {   // consumed five barriers later, rewritten 32 steps later
                const double B1 = b1t * p1, B2 = b2t * p2;
                S.sctab[2 * lane] = (float)((double)a.lr / (1.0 - B1));
                S.sctab[2 * lane + 1] = (float)(1.0 / sqrt(1.0 - B2));
                b1t = __shfl_sync(0xffffffffu, B1, 31);
                b2t = __shfl_sync(0xffffffffu, B2, 31);
            }
            // ---- forward ----
            if constexpr (MMA) {
                // TF32 remainders of the batch: computed one step ahead (see the output layer); only the first batch of the
                // launch is split here, one float4 per thread (16 rows x 128 columns)
                if (q == 0) {
                    const int o = (tid >> 5) * XS + (tid & 31) * 4;
                    const float4 xv = *reinterpret_cast<const float4*>(X + o);
                    *reinterpret_cast<float4*>(S.xlo[buf] + o) = make_float4(tf32_lo(xv.x), tf32_lo(xv.y), tf32_lo(xv.z), tf32_lo(xv.w));
                    __syncthreads();
                }
                FM_TICK(11);
                // h1[16 x 32] = X[16 x 120] W1^T: warp (kq, nh) = (k-steps kq, kq+4, .., output columns 16nh..16nh+15)
                if (warp < 8) {
                    const int g = lane >> 2, t = lane & 3, kq = warp & 3, nh = warp >> 2;
                    Acc2 c0, c1;
                    c0.zero(); c1.zero();
                    const int aoff = ((lane & 7) + ((lane >> 3) & 1) * 8) * XS + (lane >> 4) * 4;
                    const int boff = S_W1 + (8 * (2 * nh + (lane >> 4)) + (lane & 7)) * W1P + ((lane >> 3) & 1) * 4;
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int ks = kq + 4 * i;               // K = 115 -> 15 steps (x is zero past column 114)
                        if (ks < 15) {
                            uint32_t ah[4], al[4], bh[4], bl[4];
                            ldsm4(bh, S.W + boff + 8 * ks);
                            ldsm4(bl, S.Wlo + boff + 8 * ks);
                            ldsm4(ah, X + aoff + 8 * ks);
                            ldsm4(al, S.xlo[buf] + aoff + 8 * ks);
                            mma3x2(c0, c1, ah, al, bh, bl);
                        }
                    }
                    float* pp = S.part + (kq * 16 + g) * 32 + 16 * nh + 2 * t;
                    *reinterpret_cast<float2*>(pp) = make_float2(c0.sum(0), c0.sum(1));
                    *reinterpret_cast<float2*>(pp + 8) = make_float2(c1.sum(0), c1.sum(1));
                    *reinterpret_cast<float2*>(pp + 8 * 32) = make_float2(c0.sum(2), c0.sum(3));
                    *reinterpret_cast<float2*>(pp + 8 * 32 + 8) = make_float2(c1.sum(2), c1.sum(3));
                }
            } else {
                layer_nt<BT, H, 29, 4, true>(X, XS, S.W + S_W1, W1P, S.W + S_B1, S.h1, SM::HP);
            }
            FM_TICK_WORK(8);
            __syncthreads();
            FM_TICK(1);
            // bottleneck 27 -> 7 -> 27: one warp per batch row, no block barrier between the two layers
            for (int b = warp; b < BT; b += NT / 32) {
                if constexpr (MMA) {       // layer-1 epilogue: add the k-quarter partial sums, bias, ReLU
                    if (lane < H) {
                        const float* pp = S.part + b * 32 + lane;
                        S.h1[b * SM::HP + lane] = fmaxf(((pp[0] + pp[512]) + (pp[1024] + pp[1536])) + S.W[S_B1 + lane], 0.f);
                    }
                    __syncwarp();
                }
                const int i = lane >> 2, sK = lane & 3;
                float zp = 0.f;
                if (i < Z) {
#pragma unroll
                    for (int j = sK; j < H; j += 4) zp = fmaf(S.h1[b * SM::HP + j], S.W[S_W2 + i * W2P + j], zp);
                }
                zp += __shfl_xor_sync(0xffffffffu, zp, 1);
                zp += __shfl_xor_sync(0xffffffffu, zp, 2);
                if (i < Z) zp += S.W[S_B2 + i];
                float zz[Z];
                float s2 = 0.f;
#pragma unroll
                for (int k = 0; k < Z; ++k) { zz[k] = __shfl_sync(0xffffffffu, zp, k * 4); s2 = fmaf(zz[k], zz[k], s2); }
                const float nz = sqrtf(s2);
                if (lane < H) {
                    const float4 wa = *reinterpret_cast<const float4*>(S.W + S_W3 + lane * W3P);
                    const float4 wb = *reinterpret_cast<const float4*>(S.W + S_W3 + lane * W3P + 4);
                    float acc = S.W[S_B3 + lane];
                    acc = fmaf(zz[0], wa.x, fmaf(zz[1], wa.y, fmaf(zz[2], wa.z, fmaf(zz[3], wa.w, acc))));
                    acc = fmaf(zz[4], wb.x, fmaf(zz[5], wb.y, fmaf(zz[6], wb.z, acc)));
                    const float hv = fmaxf(acc, 0.f);
                    S.h2[b * SM::HP + lane] = hv;
                    if (MMA) S.h2lo[b * SM::HP + lane] = tf32_lo(hv);
                }
                const float zmine = __shfl_sync(0xffffffffu, zp, (lane < Z ? lane : 0) * 4);
                if (lane < Z) S.z[b * SM::ZP + lane] = zmine;
                if (lane == 0) {
                    S.znorm[b] = nz;
                    if (b < rows && a.shrink_lambda != 0.f) {
                        const float contrib = a.shrink_lambda * nz / (float)rows;
                        if (is_train) loss_acc += contrib; else vloss_acc += contrib;
                    }
                }
            }
            __syncthreads();
            FM_TICK(2);
            // output layer fused with the reconstruction loss and its gradient
            if constexpr (MMA) {
                // xhat[16 x 120] = h2[16 x 32] W4^T: warp w owns features 16w..16w+15 (features >= 115 are masked)
                if (warp < 8) {
                    const int g = lane >> 2, t = lane & 3;
                    Acc2 c0, c1;
                    c0.zero(); c1.zero();
                    const int aoff = ((lane & 7) + ((lane >> 3) & 1) * 8) * SM::HP + (lane >> 4) * 4;
                    const int btile = min(2 * warp + (lane >> 4), 14);      // the 16th tile does not exist: re-read the 15th
                    const int boff = S_W4 + (8 * btile + (lane & 7)) * W4P + ((lane >> 3) & 1) * 4;
#pragma unroll
                    for (int ks = 0; ks < 4; ++ks) {                        // h2 is zero past column 26
                        uint32_t ah[4], al[4], bh[4], bl[4];
                        ldsm4(bh, S.W + boff + 8 * ks);
                        ldsm4(bl, S.Wlo + boff + 8 * ks);
                        ldsm4(ah, S.h2 + aoff + 8 * ks);
                        ldsm4(al, S.h2lo + aoff + 8 * ks);
                        mma3x2(c0, c1, ah, al, bh, bl);
                    }
                    const float inv = 1.f / ((float)rows * (float)D);
                    const float sc2 = 2.f * inv;
                    float lsum = 0.f;
#pragma unroll
                    for (int nt = 0; nt < 2; ++nt) {
                        const Acc2& c = nt ? c1 : c0;
                        const int d0 = 16 * warp + 8 * nt + 2 * t;
                        if (d0 < 120) {
                            const float bb0 = S.W[S_B4 + d0], bb1 = S.W[S_B4 + d0 + 1];
#pragma unroll
                            for (int half = 0; half < 2; ++half) {
                                const int b = g + 8 * half;
                                const float2 xv = *reinterpret_cast<const float2*>(X + b * XS + d0);
                                float e0 = (c.sum(2 * half) + bb0) - xv.x;
                                float e1 = (c.sum(2 * half + 1) + bb1) - xv.y;
                                e0 = (b < rows && d0 < D) ? e0 : 0.f;
                                e1 = (b < rows && d0 + 1 < D) ? e1 : 0.f;
                                lsum = fmaf(e0, e0, fmaf(e1, e1, lsum));
                                if (is_train) {
                                    const float s0 = e0 * sc2, s1 = e1 * sc2;
                                    *reinterpret_cast<float2*>(S.dxr + b * 132 + d0) = make_float2(s0, s1);
                                    *reinterpret_cast<float2*>(S.dxrlo + b * 132 + d0) = make_float2(tf32_lo(s0), tf32_lo(s1));
                                    if (b < BT) {
                                        if (d0 < D) S.dxT[d0 * BT + b] = s0;
                                        if (d0 + 1 < D) S.dxT[(d0 + 1) * BT + b] = s1;
                                    }
                                }
                            }
                        }
                    }
                    if (is_train) loss_acc = fmaf(lsum, inv, loss_acc); else vloss_acc = fmaf(lsum, inv, vloss_acc);
                } else if (has_next) {
                    // idle here: split the NEXT batch (requested at the top of this step, long since landed) into its TF32
                    // remainders so that the next step starts without a pass + block barrier of its own
                    const int nb = buf ^ 1;
                    mbar_wait(&S.bar[nb], (nb ? uses1 : uses0) & 1u);      // the wait every thread repeats at the next step
                    const float* Xn = &S.x[nb][0];
                    float* Ln = S.xlo[nb];
#pragma unroll
                    for (int h2i = 0; h2i < 2; ++h2i) {
                        const int e = (tid - 256) + 256 * h2i;                // 512 float4: 16 rows x 32
                        const int o = (e >> 5) * XS + (e & 31) * 4;
                        const float4 xv = *reinterpret_cast<const float4*>(Xn + o);
                        *reinterpret_cast<float4*>(Ln + o) = make_float4(tf32_lo(xv.x), tf32_lo(xv.y), tf32_lo(xv.z), tf32_lo(xv.w));
                    }
                }
            } else {
                constexpr int ITEMS = D * (BT / 4);
                const float inv = 1.f / ((float)rows * (float)D);
                float lsum = 0.f;
                for (int it = tid; it < ITEMS; it += NT) {
                    const int d = it % D, bg = it / D;
                    const float* wrow = S.W + S_W4 + d * W4P;
                    const float* arow = S.h2 + (bg * 4) * SM::HP;
                    float acc0 = 0.f, acc1 = 0.f, acc2 = 0.f, acc3 = 0.f;
#pragma unroll
                    for (int cc = 0; cc < 7; ++cc) {
                        const float4 w = *reinterpret_cast<const float4*>(wrow + 4 * cc);
                        const float4 a0 = *reinterpret_cast<const float4*>(arow + 4 * cc);
                        const float4 a1 = *reinterpret_cast<const float4*>(arow + SM::HP + 4 * cc);
                        const float4 a2 = *reinterpret_cast<const float4*>(arow + 2 * SM::HP + 4 * cc);
                        const float4 a3 = *reinterpret_cast<const float4*>(arow + 3 * SM::HP + 4 * cc);
                        acc0 = fmaf(a0.x, w.x, fmaf(a0.y, w.y, fmaf(a0.z, w.z, fmaf(a0.w, w.w, acc0))));
                        acc1 = fmaf(a1.x, w.x, fmaf(a1.y, w.y, fmaf(a1.z, w.z, fmaf(a1.w, w.w, acc1))));
                        acc2 = fmaf(a2.x, w.x, fmaf(a2.y, w.y, fmaf(a2.z, w.z, fmaf(a2.w, w.w, acc2))));
                        acc3 = fmaf(a3.x, w.x, fmaf(a3.y, w.y, fmaf(a3.z, w.z, fmaf(a3.w, w.w, acc3))));
                    }
                    const float bb = S.W[S_B4 + d];
                    const int b0 = bg * 4;
                    float e0 = (acc0 + bb) - X[(b0 + 0) * XS + d];
                    float e1 = (acc1 + bb) - X[(b0 + 1) * XS + d];
                    float e2 = (acc2 + bb) - X[(b0 + 2) * XS + d];
                    float e3 = (acc3 + bb) - X[(b0 + 3) * XS + d];
                    e0 = (b0 + 0 < rows) ? e0 : 0.f; e1 = (b0 + 1 < rows) ? e1 : 0.f;
                    e2 = (b0 + 2 < rows) ? e2 : 0.f; e3 = (b0 + 3 < rows) ? e3 : 0.f;
                    lsum = fmaf(e0, e0, fmaf(e1, e1, fmaf(e2, e2, fmaf(e3, e3, lsum))));
                    if (is_train) {
                        const float sc2 = 2.f * inv;
                        *reinterpret_cast<float4*>(S.dxT + d * BT + b0) = make_float4(e0 * sc2, e1 * sc2, e2 * sc2, e3 * sc2);
                    }
                }
                if (is_train) loss_acc = fmaf(lsum, inv, loss_acc); else vloss_acc = fmaf(lsum, inv, vloss_acc);
            }
            FM_TICK_WORK(9);
            __syncthreads();
            FM_TICK(3);
            if (!is_train) continue;
            // ---- backward ----
            if constexpr (MMA) {
                // dp2[16 x 32] = relu'(h2) * (dx[16 x 120] W4[120 x 32]): warp (kq, nh) as in layer 1; the mask and the sum
                // over the k-quarters are applied by the bottleneck backward below
                if (warp < 8) {
                    const int g = lane >> 2, t = lane & 3, kq = warp & 3, nh = warp >> 2;
                    Acc2 c0, c1;
                    c0.zero(); c1.zero();
                    const int aoff = ((lane & 7) + ((lane >> 3) & 1) * 8) * 132 + (lane >> 4) * 4;
                    const int boff = S_W4 + t * W4P + 16 * nh + g;                 // B[k = d][n = j] = W4[d][j]
#pragma unroll
                    for (int i = 0; i < 4; ++i) {
                        const int ks = kq + 4 * i;                                 // dx is zero past feature 114
                        if (ks < 15) {
                            uint32_t ah[4], al[4], bh[4], bl[4];
                            const int o0 = boff + (8 * ks) * W4P, o1 = o0 + 4 * W4P;
                            bh[0] = __float_as_uint(S.W[o0]); bh[1] = __float_as_uint(S.W[o1]);
                            bh[2] = __float_as_uint(S.W[o0 + 8]); bh[3] = __float_as_uint(S.W[o1 + 8]);
                            bl[0] = __float_as_uint(S.Wlo[o0]); bl[1] = __float_as_uint(S.Wlo[o1]);
                            bl[2] = __float_as_uint(S.Wlo[o0 + 8]); bl[3] = __float_as_uint(S.Wlo[o1 + 8]);
                            ldsm4(ah, S.dxr + aoff + 8 * ks);
                            ldsm4(al, S.dxrlo + aoff + 8 * ks);
                            mma3x2(c0, c1, ah, al, bh, bl);
                        }
                    }
                    float* pp = S.part + (kq * 16 + g) * 32 + 16 * nh + 2 * t;
                    *reinterpret_cast<float2*>(pp) = make_float2(c0.sum(0), c0.sum(1));
                    *reinterpret_cast<float2*>(pp + 8) = make_float2(c1.sum(0), c1.sum(1));
                    *reinterpret_cast<float2*>(pp + 8 * 32) = make_float2(c0.sum(2), c0.sum(3));
                    *reinterpret_cast<float2*>(pp + 8 * 32 + 8) = make_float2(c1.sum(2), c1.sum(3));
                } else if (tid - 256 < D) {
                    // idle in this phase: the output-layer bias gradient is complete (dxT) and nobody reads b4 before the
                    // next step, so its Adam update leaves the two "small" warps that bound the update phase (one bias
                    // per thread on warps 8-11: short enough to hide behind the MMAs of warps 0-7)
                    AdamStep st;
                    st.step_size = S.sctab[2 * (ts & 31)]; st.inv_sqrt_bc2 = S.sctab[2 * (ts & 31) + 1]; st.mu = a.fedprox_mu;
                    float prox_b4 = 0.f;
                    const int n = tid - 256;
                    adam_update(S.W, S.M, S.V, As, S_B4 + n, row_sum<BT>(S.dxT + n * BT), st, prox_b4);
                    if (prox) loss_acc = fmaf(a.fedprox_mu, prox_b4, loss_acc);
                }
            } else {
                layer_dgrad<BT, D, H, 4, true>(S.dxT, S.W + S_W4, W4P, S.h2, SM::HP, S.dp2T);
            }
            FM_TICK_WORK(10);
            __syncthreads();
            FM_TICK(4);
            // bottleneck backward, one warp per batch row: dz = dp2 W3 + shrink gradient, dp1 = relu'(h1) * (dz W2)
            for (int b = warp; b < BT; b += NT / 32) {
                if constexpr (MMA) {       // data-gradient epilogue: add the k-quarter partial sums, ReLU mask
                    if (lane < H) {
                        const float* pp = S.part + b * 32 + lane;
                        const float gsum = (pp[0] + pp[512]) + (pp[1024] + pp[1536]);
                        S.dp2T[lane * BT + b] = S.h2[b * SM::HP + lane] > 0.f ? gsum : 0.f;
                    }
                    __syncwarp();
                }
                const int i = lane >> 2, sK = lane & 3;
                float dzp = 0.f;
                if (i < Z) {
#pragma unroll
                    for (int j = sK; j < H; j += 4) dzp = fmaf(S.dp2T[j * BT + b], S.W[S_W3 + j * W3P + i], dzp);
                }
                dzp += __shfl_xor_sync(0xffffffffu, dzp, 1);
                dzp += __shfl_xor_sync(0xffffffffu, dzp, 2);
                if (i < Z && a.shrink_lambda != 0.f && b < rows) {
                    const float nz = S.znorm[b];
                    if (nz > 0.f) dzp += a.shrink_lambda / (float)rows * S.z[b * SM::ZP + i] / nz;
                }
                float dzz[Z];
#pragma unroll
                for (int k = 0; k < Z; ++k) dzz[k] = __shfl_sync(0xffffffffu, dzp, k * 4);
                const float dzmine = __shfl_sync(0xffffffffu, dzp, (lane < Z ? lane : 0) * 4);
                if (lane < Z) S.dzT[lane * BT + b] = dzmine;
                if (lane < H) {
                    float gsum = 0.f;
#pragma unroll
                    for (int k = 0; k < Z; ++k) gsum = fmaf(dzz[k], S.W[S_W2 + k * W2P + lane], gsum);
                    S.dp1T[lane * BT + b] = S.h1[b * SM::HP + lane] > 0.f ? gsum : 0.f;
                }
            }
            __syncthreads();
            FM_TICK(5);
            // ---- all parameter updates in one balanced phase (roles are warp aligned) ----
            {
                float prox_acc = 0.f;
                AdamStep st;
                st.step_size = S.sctab[2 * (ts & 31)]; st.inv_sqrt_bc2 = S.sctab[2 * (ts & 31) + 1]; st.mu = a.fedprox_mu;
                if (tid < ROLE_W4) {
                    if (tid < W1_TILES)
                        wgrad_adam_tile<BT, H, 29>(tid, S.dp1T, X, XS, S_W1, W1P, S.W, MMA ? S.Wlo : nullptr, As, st, tm, tv, prox_acc);
                } else if (tid < ROLE_SMALL) {
                    if (tid - ROLE_W4 < W4_TILES)
                        wgrad_adam_tile<BT, D, 7>(tid - ROLE_W4, S.dxT, S.h2, SM::HP, S_W4, W4P, S.W, MMA ? S.Wlo : nullptr, As, st, tm, tv, prox_acc);
                } else {
                    static_assert(NT - ROLE_SMALL == 64 && H * 2 + 7 <= 64 && Z * 7 + 2 <= 64 && H + D <= 3 * 64, "small-role item map");
                    const int t0 = tid - ROLE_SMALL;
                    // -- gradients of this thread's W3 (+ b3) item and W2 (+ b2) item
                    float4 gA = make_float4(0.f, 0.f, 0.f, 0.f), gB = gA;
                    int iA = -1, iB = -1;
                    if (t0 < H * 2) { const int n = t0 >> 1, kc = t0 & 1; gA = wgrad_quad<BT>(S.dp2T + n * BT, S.z + 4 * kc, SM::ZP); iA = S_W3 + n * W3P + 4 * kc; }
                    else if (t0 < H * 2 + 7) { const int n0 = (t0 - H * 2) * 4; gA = bias_quad<BT, H>(S.dp2T, n0); iA = S_B3 + n0; }
                    if (t0 < Z * 7) { const int n = t0 / 7, kc = t0 - 7 * n; gB = wgrad_quad<BT>(S.dzT + n * BT, S.h1 + 4 * kc, SM::HP); iB = S_W2 + n * W2P + 4 * kc; }
                    else if (t0 < Z * 7 + 2) { const int n0 = (t0 - Z * 7) * 4; gB = bias_quad<BT, Z>(S.dzT, n0); iB = S_B2 + n0; }
                    // -- loads, Adam, stores (distinct addresses; grouped so that no load waits behind a store)
                    const float4 z4 = make_float4(0.f, 0.f, 0.f, 0.f);
                    const int jA = iA < 0 ? 0 : iA, jB = iB < 0 ? 0 : iB;
                    float4 wA = *reinterpret_cast<float4*>(S.W + jA), mA = *reinterpret_cast<float4*>(S.M + jA), vA = *reinterpret_cast<float4*>(S.V + jA);
                    float4 wB = *reinterpret_cast<float4*>(S.W + jB), mB = *reinterpret_cast<float4*>(S.M + jB), vB = *reinterpret_cast<float4*>(S.V + jB);
                    float4 aA = z4, aB = z4;
                    if (prox) { aA = *reinterpret_cast<const float4*>(As + jA); aB = *reinterpret_cast<const float4*>(As + jB); }
                    if (iA >= 0) {
                        adam1(wA.x, mA.x, vA.x, aA.x, gA.x, st, prox_acc); adam1(wA.y, mA.y, vA.y, aA.y, gA.y, st, prox_acc);
                        adam1(wA.z, mA.z, vA.z, aA.z, gA.z, st, prox_acc); adam1(wA.w, mA.w, vA.w, aA.w, gA.w, st, prox_acc);
                    }
                    if (iB >= 0) {
                        adam1(wB.x, mB.x, vB.x, aB.x, gB.x, st, prox_acc); adam1(wB.y, mB.y, vB.y, aB.y, gB.y, st, prox_acc);
                        adam1(wB.z, mB.z, vB.z, aB.z, gB.z, st, prox_acc); adam1(wB.w, mB.w, vB.w, aB.w, gB.w, st, prox_acc);
                    }
                    if (iA >= 0) { *reinterpret_cast<float4*>(S.W + iA) = wA; *reinterpret_cast<float4*>(S.M + iA) = mA; *reinterpret_cast<float4*>(S.V + iA) = vA; }
                    if (iB >= 0) { *reinterpret_cast<float4*>(S.W + iB) = wB; *reinterpret_cast<float4*>(S.M + iB) = mB; *reinterpret_cast<float4*>(S.V + iB) = vB; }
                    // -- biases of the two big layers as a second, short stage (keeps the register peak down); with the
                    //    tensor-core variant b4 was already updated by idle warps of the data-gradient phase
                    constexpr int NC = MMA ? 1 : 3;
                    int iC[3];
                    float gC[3], wC[3], mC[3], vC[3], aC[3];
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        const int n = t0 + 64 * i;
                        iC[i] = -1; gC[i] = 0.f;
                        if (i < NC) {
                            if (n < H) { gC[i] = row_sum<BT>(S.dp1T + n * BT); iC[i] = S_B1 + n; }
                            else if (!MMA && n < H + D) { gC[i] = row_sum<BT>(S.dxT + (n - H) * BT); iC[i] = S_B4 + (n - H); }
                        }
                    }
#pragma unroll
                    for (int i = 0; i < 3; ++i) {
                        const int j = iC[i] < 0 ? 0 : iC[i];
                        wC[i] = S.W[j]; mC[i] = S.M[j]; vC[i] = S.V[j]; aC[i] = prox ? As[j] : 0.f;
                    }
#pragma unroll
                    for (int i = 0; i < 3; ++i) if (iC[i] >= 0) adam1(wC[i], mC[i], vC[i], aC[i], gC[i], st, prox_acc);
#pragma unroll
                    for (int i = 0; i < 3; ++i) if (iC[i] >= 0) { S.W[iC[i]] = wC[i]; S.M[iC[i]] = mC[i]; S.V[iC[i]] = vC[i]; }
                }
                if (prox) loss_acc = fmaf(a.fedprox_mu, prox_acc, loss_acc);
            }
            ++ts;
            __syncthreads();
            FM_TICK(6);
        }
        // ---- epoch bookkeeping (reference: client_trainer.py:385, :404-417) ----
        float prox_val = 0.f;
        if (prox) {
            float pa = 0.f;
            for (int p = tid; p < NP; p += NT) { const int s = sidx(p); const float dl = S.W[s] - As[s]; pa = fmaf(dl, dl, pa); }
            prox_val = a.fedprox_mu * (float)block_sum<double>((double)pa, S.red);
        }
        const double tl = block_sum<double>((double)loss_acc, S.red);
        const double vl = block_sum<double>((double)vloss_acc, S.red);
        const float train_loss = (float)(tl / (double)max(nbt, 1));
        const float valid_loss = nbv > 0 ? (float)(vl / (double)nbv) + prox_val : __int_as_float(0x7fc00000);
        if (tid == 0 && e < a.log_stride) {
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 0] = train_loss;
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 1] = valid_loss;
        }
        epochs_done = e + 1;
        bool stop = false;
        if (nbv == 0) {
            // no validation rows = the reference's run(train_loader, valid_loader=None): all epochs, no early
            // stop, no best-epoch snapshot (client_trainer.py:387 guards the whole block with `if valid_loader`)
        } else if (valid_loss < min_valid) {      // uniform across the block: every thread holds the same value
            min_valid = valid_loss; worse = 0;
            float* gb = a.best + (size_t)c * P_PAD;
            for (int p = tid; p < NP; p += NT) gb[p] = S.W[sidx(p)];
        } else if (++worse >= a.patience) {
            stop = true;
        }
        if (stop) break;
    }
    // drain a speculative prefetch so no async write lands after the CTA exits
    if (outstanding) { const int buf = (int)(q & 1); mbar_wait(&S.bar[buf], (buf ? uses1 : uses0) & 1u); }
    __syncthreads();
    for (int p = tid; p < NP; p += NT) {
        const int s = sidx(p);
        gp[p] = S.W[s];
        const int tt = tensor_of(p);
        if (tt != 0 && tt != 6) { gm[p] = S.M[s]; gv[p] = S.V[s]; }     // W1 / W4 moments live in registers
    }
    if (tid < ROLE_W4) { if (tid < W1_TILES) tile_moments_io<H, D, 29>(tid, OFF_W1, gm, gv, tm, tv, true); }
    else if (tid < ROLE_SMALL) { if (tid - ROLE_W4 < W4_TILES) tile_moments_io<D, H, 7>(tid - ROLE_W4, OFF_W4, gm, gv, tm, tv, true); }
    if (dbg) { FM_TICK(7); for (int i = 0; i < 16; ++i) a.phase_cycles[i] = (unsigned long long)dbg_acc[i] + (dbg_sink & 0); }
    if (tid == NT - 1) a.adam_t[c] = t_step0 + ts;
    if (tid == 0) {
        a.epochs_run[c] = epochs_done;
        a.trained_rounds[c] += 1;
        for (int e = epochs_done; e < a.log_stride; ++e) {
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 0] = 0.f;
            a.loss_log[((size_t)c * a.log_stride + e) * 2 + 1] = 0.f;
        }
    }
}

}  // namespace fm

// ---------------------------------------------------------------------------------------------
extern "C" int fm_train_clients(const FmTrainArgs* args, void* stream) {
    using namespace fm;
    if (args->n_sel <= 0) return 0;
    if (args->batch <= 0 || args->batch > 32) return (int)cudaErrorInvalidValue;
    const bool prox = args->fedprox_mu != 0.f;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t err;
    static int use_mma = -1;
    if (use_mma < 0) { const char* e = getenv("FM_TRAIN_MMA"); use_mma = (e && e[0] == '0') ? 0 : 1; }
    if (args->batch <= 12 && use_mma) {
        size_t smem = ((sizeof(TrainSmem<12, true>) + 15) / 16) * 16 + (prox ? S_TOT * sizeof(float) : 0);
        err = cudaFuncSetAttribute(train_clients_kernel<12, true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return (int)err;
        train_clients_kernel<12, true><<<args->n_sel, NT, smem, st>>>(*args);
    } else if (args->batch <= 12) {
        size_t smem = ((sizeof(TrainSmem<12, false>) + 15) / 16) * 16 + (prox ? S_TOT * sizeof(float) : 0);
        err = cudaFuncSetAttribute(train_clients_kernel<12, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return (int)err;
        train_clients_kernel<12, false><<<args->n_sel, NT, smem, st>>>(*args);
    } else {
        size_t smem = ((sizeof(TrainSmem<32, false>) + 15) / 16) * 16 + (prox ? S_TOT * sizeof(float) : 0);
        err = cudaFuncSetAttribute(train_clients_kernel<32, false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return (int)err;
        train_clients_kernel<32, false><<<args->n_sel, NT, smem, st>>>(*args);
    }
    return (int)cudaGetLastError();
}
