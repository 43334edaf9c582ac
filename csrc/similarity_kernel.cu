// T6 -- the weight function of the historical `fusion_avg` merge rule on the device.
//
// Reference: src/Utils/utils.py:10-24 -- fit a Gaussian KDE (Scott bandwidth) on a data set, take its log-density at
// its own points, and return the Jensen-Shannon distance between exp(those scores) and the dev-set scores.  In this
// framework the "generated data set" of client j is the dev rows as reconstructed by client j's autoencoder
// (fedmse_decentralized_b200/utils/similarity.py:fusion_distances is the fp64 torch oracle of exactly this).
//
//   kde_logdens_kernel   one CTA per selected model (+ one for the dev rows themselves): forward of the first n rows
//                        through the model, then log p(r_i) = logsumexp_k(-|r_i - r_k|^2 / 2h^2) + const for every row,
//                        pairwise differences formed directly (no |a|^2 + |b|^2 - 2ab cancellation), online
//                        log-sum-exp, the k rows staged through shared memory 32 at a time.
//   js_distance_kernel   one CTA per model: sqrt(JS(softmax(ref), softmax(dens_j))) -- scipy's jensenshannon
//                        normalises both vectors, so only the softmax of the log-densities matters.
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

constexpr int KD_NT = 256;
constexpr int KD_TILE = 32;
constexpr int KD_TP = 116;                       // tile row pitch (floats): 16-byte aligned rows

struct KdSmem {
    float w[P_PAD];
    float tile[KD_TILE][KD_TP];
};

__global__ void __launch_bounds__(KD_NT) kde_logdens_kernel(const FmKdeJsArgs a) {
    extern __shared__ __align__(16) unsigned char kd_raw[];
    KdSmem& S = *reinterpret_cast<KdSmem*>(kd_raw);
    const int tid = threadIdx.x, slot = blockIdx.x, n = a.n_rows;
    const bool is_ref = slot == a.n_slots;
    const float* rec = a.X;                      // reference CTA: the rows themselves (pitch XP)
    int pitch = XP;
    if (!is_ref) {
        const int c = (a.slot_client ? a.slot_client[slot] : a.first_client + slot) - a.first_client;
        if (c < 0 || c >= a.n_local) return;     // not this rank's model
        const float* row = a.slab + (size_t)c * P_PAD;
        for (int i = tid; i < NP; i += KD_NT) S.w[i] = row[i];
        __syncthreads();
        float* out = a.rec_scratch + (size_t)slot * n * XP;
        for (int i = tid; i < n; i += KD_NT) {   // one row per thread: 115 -> 27 -> 7 -> 27 -> 115
            const float* x = a.X + (size_t)i * XP;
            float h1[H], z[Z], h2[H];
#pragma unroll
            for (int j = 0; j < H; ++j) h1[j] = S.w[OFF_B1 + j];
            for (int k = 0; k < D; ++k) {
                const float xv = x[k];
#pragma unroll
                for (int j = 0; j < H; ++j) h1[j] = fmaf(S.w[OFF_W1 + j * D + k], xv, h1[j]);
            }
#pragma unroll
            for (int j = 0; j < H; ++j) h1[j] = fmaxf(h1[j], 0.f);
#pragma unroll
            for (int q = 0; q < Z; ++q) {
                float acc = S.w[OFF_B2 + q];
#pragma unroll
                for (int j = 0; j < H; ++j) acc = fmaf(S.w[OFF_W2 + q * H + j], h1[j], acc);
                z[q] = acc;
            }
#pragma unroll
            for (int j = 0; j < H; ++j) {
                float acc = S.w[OFF_B3 + j];
#pragma unroll
                for (int q = 0; q < Z; ++q) acc = fmaf(S.w[OFF_W3 + j * Z + q], z[q], acc);
                h2[j] = fmaxf(acc, 0.f);
            }
            for (int d = 0; d < D; ++d) {
                float acc = S.w[OFF_B4 + d];
#pragma unroll
                for (int j = 0; j < H; ++j) acc = fmaf(S.w[OFF_W4 + d * H + j], h2[j], acc);
                out[(size_t)i * XP + d] = acc;
            }
        }
        __syncthreads();                         // the reconstructions of this CTA are visible to all of its threads
        rec = out;
    }
    // Scott bandwidth (sklearn): h = n^(-1 / (d + 4)); log of the Gaussian normalisation and of 1 / n
    const double hd = pow((double)n, -1.0 / (double)(D + 4));
    const float neg_half_inv_h2 = (float)(-0.5 / (hd * hd));
    const float log_norm = (float)(-0.5 * D * log(2.0 * 3.14159265358979323846) - D * log(hd) - log((double)n));
    for (int i0 = 0; i0 < n; i0 += KD_NT) {
        const int i = i0 + tid;
        const bool live = i < n;
        float xi[D];
#pragma unroll
        for (int c = 0; c < D; ++c) xi[c] = live ? rec[(size_t)i * pitch + c] : 0.f;
        float m = -INFINITY, s = 0.f;
        for (int k0 = 0; k0 < n; k0 += KD_TILE) {
            __syncthreads();
            for (int e = tid; e < KD_TILE * D; e += KD_NT) {
                const int kk = e / D, c = e - kk * D;
                S.tile[kk][c] = k0 + kk < n ? rec[(size_t)(k0 + kk) * pitch + c] : 0.f;
            }
            __syncthreads();
            const int kn = min(KD_TILE, n - k0);
            for (int kk = 0; kk < kn; ++kk) {
                float d0 = 0.f, d1 = 0.f, d2 = 0.f, d3 = 0.f;
#pragma unroll
                for (int c4 = 0; c4 < D / 4; ++c4) {
                    const float4 t = *reinterpret_cast<const float4*>(&S.tile[kk][4 * c4]);
                    const float e0 = xi[4 * c4] - t.x, e1 = xi[4 * c4 + 1] - t.y, e2 = xi[4 * c4 + 2] - t.z, e3 = xi[4 * c4 + 3] - t.w;
                    d0 = fmaf(e0, e0, d0); d1 = fmaf(e1, e1, d1); d2 = fmaf(e2, e2, d2); d3 = fmaf(e3, e3, d3);
                }
#pragma unroll
                for (int c = D / 4 * 4; c < D; ++c) { const float e = xi[c] - S.tile[kk][c]; d0 = fmaf(e, e, d0); }
                const float v = ((d0 + d1) + (d2 + d3)) * neg_half_inv_h2;
                if (v > m) { s = s * __expf(m - v) + 1.f; m = v; }      // exp(-inf - v) = 0 on the first term
                else s += __expf(v - m);
            }
        }
        if (live) a.dens[(size_t)slot * n + i] = m + logf(s) + log_norm;
    }
}

__device__ __forceinline__ float block_max(float v, float* scratch) {
    const int lane = threadIdx.x & 31, w = threadIdx.x >> 5, nw = (blockDim.x + 31) >> 5;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v = fmaxf(v, __shfl_xor_sync(0xffffffffu, v, o));
    __syncthreads();
    if (lane == 0) scratch[w] = v;
    __syncthreads();
    float t = lane < nw ? scratch[lane] : -INFINITY;
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) t = fmaxf(t, __shfl_xor_sync(0xffffffffu, t, o));
    return t;
}

__global__ void __launch_bounds__(KD_NT) js_distance_kernel(const FmKdeJsArgs a) {
    __shared__ float fred[33];
    __shared__ double dred[34];
    const int tid = threadIdx.x, slot = blockIdx.x, n = a.n_rows;
    const int c = (a.slot_client ? a.slot_client[slot] : a.first_client + slot) - a.first_client;
    if (c < 0 || c >= a.n_local) return;
    const float* lp = a.dens + (size_t)a.n_slots * n;      // reference log-densities
    const float* lq = a.dens + (size_t)slot * n;
    float mp = -INFINITY, mq = -INFINITY;
    for (int i = tid; i < n; i += KD_NT) { mp = fmaxf(mp, lp[i]); mq = fmaxf(mq, lq[i]); }
    mp = block_max(mp, fred);
    mq = block_max(mq, fred);
    double sp = 0.0, sq = 0.0;
    for (int i = tid; i < n; i += KD_NT) { sp += exp((double)(lp[i] - mp)); sq += exp((double)(lq[i] - mq)); }
    sp = block_sum<double>(sp, dred);
    sq = block_sum<double>(sq, dred);
    double js = 0.0;
    for (int i = tid; i < n; i += KD_NT) {
        const double p = exp((double)(lp[i] - mp)) / sp, q = exp((double)(lq[i] - mq)) / sq, mm = 0.5 * (p + q);
        if (p > 0.0) js += 0.5 * p * (log(p) - log(mm));
        if (q > 0.0) js += 0.5 * q * (log(q) - log(mm));
    }
    js = block_sum<double>(js, dred);
    if (tid == 0) a.dist_out[slot] = (float)sqrt(js > 0.0 ? js : 0.0) + 1e-12f;
}

}  // namespace fm

extern "C" int64_t fm_kde_js_scratch_floats(int32_t n_slots, int32_t n_rows) {
    return (int64_t)(n_slots > 0 ? n_slots : 1) * n_rows * fm::XP;
}

extern "C" int fm_kde_js(const FmKdeJsArgs* args, void* stream) {
    using namespace fm;
    if (args->n_slots <= 0 || args->n_rows <= 1) return (int)cudaErrorInvalidValue;
    if (!args->X || !args->slab || !args->rec_scratch || !args->dens || !args->dist_out) return (int)cudaErrorInvalidValue;
    cudaStream_t st = (cudaStream_t)stream;
    cudaError_t err = cudaFuncSetAttribute((const void*)kde_logdens_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize,
                                           (int)sizeof(KdSmem));
    if (err != cudaSuccess) return (int)err;
    kde_logdens_kernel<<<args->n_slots + 1, KD_NT, sizeof(KdSmem), st>>>(*args);
    js_distance_kernel<<<args->n_slots, KD_NT, 0, st>>>(*args);
    return (int)cudaGetLastError();
}
