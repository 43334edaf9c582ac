// Gossip merge -- the per-client form of T2: every selected client merges the models of its topology
// neighbours (and its own), weighted by each model's reconstruction MSE on the client's *own*
// validation rows, then gates the result through its verifier.  This is the decentralised path the
// reference only sketches as an API (broadcast_model / receive_model / request_aggregation,
// src/Trainer/client_trainer.py:141-172; never called by src/main.py).
//
// One CTA per merging client fuses: neighbour model fetch (peer-mapped loads, NVLink for remote
// owners) -> scoring forward -> weight normalisation -> weighted parameter reduction -> verification
// forward + decision + update.  Models are read from a snapshot slab (synchronous gossip).
#include "score_core.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

constexpr int MAX_CAND = 1024;

struct GossipSmem {
    ScoreSmem sc;
    alignas(16) float theta[P_PAD];
    float wts[MAX_CAND];
    float wsse[8];
    double red[34];
};

__device__ __forceinline__ void owner_of_g(int g, int n_global, int world, int& rank, int& local) {
    const int base = n_global / world, rem = n_global - base * world;
    const int cut = rem * (base + 1);
    if (g < cut) { rank = g / (base + 1); local = g - rank * (base + 1); }
    else { const int q = (g - cut) / base; rank = rem + q; local = (g - cut) - q * base; }
}

// mean squared reconstruction error of the model currently loaded in S.sc over rows V[0..nr)
__device__ float model_mse(GossipSmem& S, const float* __restrict__ V, int nr) {
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    float acc = 0.f;
    for (int r0 = 0; r0 < nr; r0 += TILE) {
        const int rows = min(TILE, nr - r0);
        __syncthreads();
        for (int i = tid; i < rows * (XP / 4); i += SNT)
            reinterpret_cast<float4*>(&S.sc.x[0][0])[i] = reinterpret_cast<const float4*>(V + (size_t)r0 * XP)[i];
        __syncthreads();
        for (int grp = 0; grp < 2; ++grp) {
            const int q0 = w * 8 + grp * 4;
            if (q0 >= rows) break;
            float sse[4], zl[4];
            score4(S.sc, &S.sc.x[0][q0 * XP], w, lane, true, sse, zl);
            if (lane == 0)
                for (int r = 0; r < 4; ++r) if (q0 + r < rows) acc += sse[r];
        }
    }
    const double tot = block_sum<double>((double)acc, S.red);
    return (float)(tot / ((double)nr * (double)D));
}


__global__ void __launch_bounds__(SNT, 1) gossip_kernel(const FmGossipArgs a) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    GossipSmem& S = *reinterpret_cast<GossipSmem*>(smem_raw);
    const int tid = threadIdx.x;
    const int c = a.sel_local[blockIdx.x];
    const int g = a.first_client + c;
    const int c0 = a.nbr_off[g], nc = min(a.nbr_off[g + 1] - c0, MAX_CAND);
    const float* V = a.x_valid + (size_t)a.valid_off[g] * XP;
    const int nr = a.valid_off[g + 1] - a.valid_off[g];
    for (int i = tid; i < 2 * TILE * XP; i += SNT) (&S.sc.x[0][0])[i] = 0.f;
    __syncthreads();
    auto model_ptr = [&](int k) -> const float* {
        int r, l;
        owner_of_g(a.nbr_idx[c0 + k], a.n_global, a.world, r, l);
        return a.peer_src[r] + (size_t)l * P_PAD;
    };
    // ---- weights ----
    if (a.mode == 1) {
        for (int k = 0; k < nc; ++k) {
            load_model_transposed(S.sc, model_ptr(k));
            const float mse = model_mse(S, V, nr);
            if (tid == 0) S.wts[k] = 1.f / mse;
        }
        __syncthreads();
        double part = 0.0;
        for (int k = tid; k < nc; k += SNT) part += (double)S.wts[k];
        const double tot = block_sum<double>(part, S.red);
        for (int k = tid; k < nc; k += SNT) S.wts[k] = (float)((double)S.wts[k] / tot);
    } else {
        for (int k = tid; k < nc; k += SNT) S.wts[k] = 1.f / (float)nc;
    }
    __syncthreads();
    // ---- weighted reduction into shared memory ----
    for (int p = tid; p < P_PAD; p += SNT) {
        float acc = 0.f;
        if (p < NP) for (int k = 0; k < nc; ++k) acc = fmaf(S.wts[k], model_ptr(k)[p], acc);
        S.theta[p] = acc;
    }
    __syncthreads();
    // ---- verification gate on the client's own validation rows ----
    load_model_transposed(S.sc, S.theta);
    const float perf = 1.f / (1.f + model_mse(S, V, nr));
    const bool first = a.hist_valid[c] == 0;
    float* hist_c = a.hist + (size_t)c * P_PAD;
    float change = 0.f;
    bool ok = true;
    if (!first) {
        float part[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int p = tid; p < NP; p += SNT) { const float dl = hist_c[p] - S.theta[p]; part[tensor_of(p)] = fmaf(dl, dl, part[tensor_of(p)]); }
#pragma unroll
        for (int t = 0; t < 8; ++t) change += sqrtf((float)block_sum<double>((double)part[t], S.red));
        ok = (change <= a.thr_change) && (perf - a.hist_perf[c] >= -a.thr_perf);
    }
    __syncthreads();
    float* params_c = a.params + (size_t)c * P_PAD;
    float* anchor_c = a.anchor + (size_t)c * P_PAD;
    for (int p = tid; p < P_PAD; p += SNT) {
        const float t = S.theta[p];
        hist_c[p] = t;
        if (ok) { params_c[p] = t; anchor_c[p] = t; }
    }
    if (tid == 0) {
        a.hist_perf[c] = perf;
        a.hist_valid[c] = 1;
        a.rejected[c] = ok ? 0 : a.rejected[c] + 1;
    }
}

}  // namespace fm

extern "C" int fm_gossip(const FmGossipArgs* args, void* stream) {
    using namespace fm;
    if (args->n_sel <= 0) return 0;
    const size_t smem = sizeof(GossipSmem);
    cudaError_t err = cudaFuncSetAttribute(gossip_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    gossip_kernel<<<args->n_sel, SNT, smem, (cudaStream_t)stream>>>(*args);
    return (int)cudaGetLastError();
}
