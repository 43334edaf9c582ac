// T4 -- anomaly scores -> tie-aware ROC-AUC, one CTA per client.
//
// Reference: src/Evaluator/evaluator.py:50-119 (per-row score, sklearn roc_curve + auc on the host) and
// src/Model/Centroid.py:15-35 (standardise by the training latents, distance to the origin).
// sklearn's trapezoid over tied thresholds equals the Mann-Whitney statistic with ties counted 1/2,
// so the AUC is computed here exactly in integers: sort (score, label) pairs in shared memory with a
// bitonic network, prefix-count the negatives, and for every tie group add
// positives * (2 * negatives_below + negatives_in_group); AUC = U2 / (2 n+ n-) in double.
//
// Capacity: up to 16,384 test rows per client (npad = next power of two, 8 B per key, <= 200 KB) the keys live
// in shared memory.  Larger test sets (raw N-BaIoT devices put hundreds of thousands of abnormal rows in the test
// set, Configuration/nba-iot-training.json) use the GLOBAL variant: keys in a caller-provided global scratch
// (L2-resident), sorted as 8,192-key shared-memory chunks + global compare-exchange passes for the wide strides.
#include <float.h>
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

constexpr int ANT = 1024;

__device__ __forceinline__ uint32_t float_key(float f) {   // order-preserving float -> uint
    if (f != f) f = 0.f;                                    // nan_to_num
    else if (f == __int_as_float(0x7f800000)) f = FLT_MAX;
    else if (f == __int_as_float(0xff800000)) f = -FLT_MAX;
    if (f == 0.f) f = 0.f;                                  // -0 == +0 must tie
    const uint32_t u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}

constexpr int AUC_CHUNK = 8192;        // keys per shared-memory chunk of the global variant (64 KB)

// compare-exchange network steps j = j_hi, j_hi/2, ..., 1 of merge size k on a shared-memory chunk whose first
// key has global index `base` (the sort direction depends on the global index)
__device__ __forceinline__ void bitonic_steps_smem(unsigned long long* sk, int cnt, int base, int k, int j_hi) {
    for (int j = j_hi; j > 0; j >>= 1) {
        for (int i = threadIdx.x; i < cnt; i += ANT) {
            const int ixj = i ^ j;
            if (ixj > i) {
                const unsigned long long x = sk[i], y = sk[ixj];
                const bool up = ((base + i) & k) == 0;
                if ((x > y) == up) { sk[i] = y; sk[ixj] = x; }
            }
        }
        __syncthreads();
    }
}

template <bool GLOBAL>
__global__ void __launch_bounds__(ANT, 1) auc_kernel(const FmAucArgs a, int npad) {
    extern __shared__ __align__(16) unsigned long long skeys[];   // [npad] (or one AUC_CHUNK when GLOBAL)
    unsigned long long* keys = GLOBAL ? a.keys_scratch + (size_t)blockIdx.x * npad : skeys;
    __shared__ double red[34];
    __shared__ int chunk_neg[ANT];
    __shared__ float mu[8], isd[8];
    const int c = blockIdx.x, tid = threadIdx.x;
    const int t0 = a.test_off[c], n = a.test_off[c + 1] - t0;
    // ---- hybrid: centroid statistics of the training latents ----
    if (a.mode == 1) {
        const int r0 = a.train_off[c], nr = a.train_off[c + 1] - r0;
        for (int i = 0; i < Z; ++i) {
            double s = 0.0;
            for (int r = tid; r < nr; r += ANT) s += (double)a.train_lat[(size_t)(r0 + r) * 8 + i];
            const double mean = block_sum<double>(s, red) / (double)nr;
            double q = 0.0;
            for (int r = tid; r < nr; r += ANT) { const double dl = (double)a.train_lat[(size_t)(r0 + r) * 8 + i] - mean; q += dl * dl; }
            const double var = block_sum<double>(q, red) / (double)nr;      // population variance
            if (tid == 0) { const double sd = sqrt(var); mu[i] = (float)mean; isd[i] = sd == 0.0 ? 1.f : (float)(1.0 / sd); }
        }
        __syncthreads();
    }
    // ---- scores -> sortable keys (label in the low word) ----
    int my_pos = 0;
    for (int r = tid; r < npad; r += ANT) {
        unsigned long long k = ~0ull;                        // padding sorts last
        if (r < n) {
            float s;
            if (a.mode == 0) s = a.test_sse[t0 + r] / (float)D;
            else {
                float q = 0.f;
#pragma unroll
                for (int i = 0; i < Z; ++i) { const float v = (a.test_lat[(size_t)(t0 + r) * 8 + i] - mu[i]) * isd[i]; q = fmaf(v, v, q); }
                s = sqrtf(q);
            }
            if (a.scores_out) a.scores_out[t0 + r] = s;
            const uint32_t lab = a.labels[t0 + r] > 0.5f ? 1u : 0u;
            my_pos += (int)lab;
            k = ((unsigned long long)float_key(s) << 32) | lab;
        }
        keys[r] = k;
    }
    const long long n_pos = (long long)block_sum<double>((double)my_pos, red);
    const long long n_neg = (long long)n - n_pos;
    __syncthreads();
    // ---- bitonic sort of npad keys ----
    if (!GLOBAL) {
        for (int k = 2; k <= npad; k <<= 1) bitonic_steps_smem(keys, npad, 0, k, k >> 1);
    } else {
        __threadfence_block();
        // (A) every chunk completely sorted in shared memory (all merge sizes up to the chunk)
        for (int b = 0; b < npad; b += AUC_CHUNK) {
            for (int i = tid; i < AUC_CHUNK; i += ANT) skeys[i] = keys[b + i];
            __syncthreads();
            for (int k = 2; k <= AUC_CHUNK; k <<= 1) bitonic_steps_smem(skeys, AUC_CHUNK, b, k, k >> 1);
            for (int i = tid; i < AUC_CHUNK; i += ANT) keys[b + i] = skeys[i];
            __syncthreads();
        }
        // (B) wider merges: strides >= AUC_CHUNK in global memory, the rest of each merge per chunk in smem
        for (int k = 2 * AUC_CHUNK; k <= npad; k <<= 1) {
            for (int j = k >> 1; j >= AUC_CHUNK; j >>= 1) {
                for (int i = tid; i < npad; i += ANT) {
                    const int ixj = i ^ j;
                    if (ixj > i) {
                        const unsigned long long x = keys[i], y = keys[ixj];
                        const bool up = (i & k) == 0;
                        if ((x > y) == up) { keys[i] = y; keys[ixj] = x; }
                    }
                }
                __syncthreads();
            }
            for (int b = 0; b < npad; b += AUC_CHUNK) {
                for (int i = tid; i < AUC_CHUNK; i += ANT) skeys[i] = keys[b + i];
                __syncthreads();
                bitonic_steps_smem(skeys, AUC_CHUNK, b, k, AUC_CHUNK >> 1);
                for (int i = tid; i < AUC_CHUNK; i += ANT) keys[b + i] = skeys[i];
                __syncthreads();
            }
        }
    }
    // ---- negatives prefix per contiguous chunk ----
    const int per = (n + ANT - 1) / ANT;
    const int b0 = min(n, tid * per), b1 = min(n, b0 + per);
    int negs = 0;
    for (int i = b0; i < b1; ++i) negs += 1 - (int)(keys[i] & 1ull);
    chunk_neg[tid] = negs;
    __syncthreads();
    // exclusive scan over ANT chunk counts (Hillis-Steele in place, double buffered through registers)
    for (int o = 1; o < ANT; o <<= 1) {
        const int v = tid >= o ? chunk_neg[tid - o] : 0;
        __syncthreads();
        chunk_neg[tid] += v;
        __syncthreads();
    }
    long long u2 = 0;
    int run_neg = chunk_neg[tid] - negs;                     // negatives strictly before b0
    for (int i = b0; i < b1; ++i) {
        const uint32_t key = (uint32_t)(keys[i] >> 32);
        const bool start = (i == 0) || ((uint32_t)(keys[i - 1] >> 32) != key);
        if (start) {                                         // walk the whole tie group (may cross chunks)
            long long gp = 0, gn = 0;
            for (int j = i; j < n && (uint32_t)(keys[j] >> 32) == key; ++j) { if (keys[j] & 1ull) ++gp; else ++gn; }
            u2 += gp * (2ll * run_neg + gn);
        }
        run_neg += 1 - (int)(keys[i] & 1ull);
    }
    const double tot = block_sum<double>((double)u2, red);   // exact: |u2| < 2^53
    if (tid == 0) a.auc[c] = (n_pos > 0 && n_neg > 0) ? tot / (2.0 * (double)n_pos * (double)n_neg)
                                                      : __longlong_as_double(0x7ff8000000000000ll);
}

}  // namespace fm

extern "C" int fm_auc(const FmAucArgs* args, void* stream) {
    using namespace fm;
    if (args->n_local <= 0) return 0;
    int npad = 2;
    while (npad < args->max_rows) npad <<= 1;
    const size_t smem = (size_t)npad * sizeof(unsigned long long);
    cudaError_t err;
    if (smem <= 200 * 1024) {                                    // <= 16,384 test rows per client 
        err = cudaFuncSetAttribute(auc_kernel<false>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (err != cudaSuccess) return (int)err;
        auc_kernel<false><<<args->n_local, ANT, smem, (cudaStream_t)stream>>>(*args, npad);
    } else {
        // keys in global scratch: the caller provides n_local * fm_auc_scratch_keys(max_rows) 8-byte words
        if (args->keys_scratch == nullptr || npad < 2 * AUC_CHUNK) return (int)cudaErrorInvalidValue;
        const size_t csm = (size_t)AUC_CHUNK * sizeof(unsigned long long);
        err = cudaFuncSetAttribute(auc_kernel<true>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)csm);
        if (err != cudaSuccess) return (int)err;
        auc_kernel<true><<<args->n_local, ANT, csm, (cudaStream_t)stream>>>(*args, npad);
    }
    return (int)cudaGetLastError();
}

// 8-byte key slots per client the global-memory variant needs (0 = the shared-memory variant serves max_rows)
extern "C" int64_t fm_auc_scratch_keys(int32_t max_rows) {
    int npad = 2;
    while (npad < max_rows) npad <<= 1;
    return (size_t)npad * sizeof(unsigned long long) <= 200 * 1024 ? 0 : (int64_t)npad;
}
