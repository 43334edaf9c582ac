// Gossip merge -- the per-client form of T2: every selected client merges the models of its topology
// neighbours (and its own), weighted by each model's reconstruction MSE on the client's *own*
// validation rows, then gates the result through its verifier.  This is the decentralised path the
// reference only sketches as an API (broadcast_model / receive_model / request_aggregation,
// src/Trainer/client_trainer.py:141-172; never called by src/main.py).
//
// One CTA per merging client fuses: neighbour model fetch -> scoring forward -> weight -> weighted
// parameter reduction -> verification forward + decision + update.
//
//  * Every neighbour row crosses NVLink (or L2) exactly ONCE per CTA: the 27 KB row is brought into shared
//    memory by the TMA unit (1-D bulk copy + mbarrier; for a remote owner the source is a peer-mapped address,
//    a plain unicast load over NVSwitch), double buffered so that row k+1 travels while row k is scored.  The
//    same shared-memory copy then feeds the scoring forward and the reduction: theta is accumulated as
//    sum_j s_j * row_j while row_j is resident and divided by sum_j s_j at the end, so no row is kept or
//    fetched twice and the neighbour count is unbounded (full mesh = every other client).
//  * Cross-GPU ordering without NCCL: models are read from a double-buffered snapshot slab.  fm_gossip_publish
//    copies the trained parameters into snapshot[round & 1] and bumps arrive[3] on every rank with a release
//    reduction; the merging CTAs spin on their own counter with acquire loads.  A buffer is rewritten two rounds
//    later, after its owner has passed the next round's wait, which every peer only reaches after finishing the
//    reads of this round -- so one counter serves both directions.
//  * The neighbour table is generated on the device: full mesh and ring are formulas of the client id, random-k
//    is a counter-based hash of (seed, round, client, draw) that parallel/topology.py restates bit for bit for
//    the oracle; an explicit CSR table is still accepted for custom graphs.
#include "score_core.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

constexpr int GOSSIP_MAX_K = 64;          // random-k / ring neighbour list kept in shared memory

struct GossipSmem {
    ScoreSmem sc;
    alignas(128) float raw[2][P_PAD];     // neighbour rows in flight / in use
    alignas(16) float theta[P_PAD];       // running sum_j s_j * row_j, then the merged model
    int nbr[GOSSIP_MAX_K + 1];
    double red[34];
    alignas(8) uint64_t row_bar[2];
};

__device__ __forceinline__ void owner_of_g(int g, int n_global, int world, int& rank, int& local) {
    const int base = n_global / world, rem = n_global - base * world;
    const int cut = rem * (base + 1);
    if (g < cut) { rank = g / (base + 1); local = g - rank * (base + 1); }
    else { const int q = (g - cut) / base; rank = rem + q; local = (g - cut) - q * base; }
}

__host__ __device__ __forceinline__ unsigned long long splitmix64(unsigned long long x) {
    x += 0x9E3779B97F4A7C15ull;
    x = (x ^ (x >> 30)) * 0xBF58476D1CE4E5B9ull;
    x = (x ^ (x >> 27)) * 0x94D049BB133111EBull;
    return x ^ (x >> 31);
}
// draw `attempt` of neighbour slot `j` of client g in round r: uniform over the n-1 other clients
__host__ __device__ __forceinline__ int random_neighbour(unsigned long long seed, int round, int g, int j, int attempt, int n) {
    unsigned long long h = splitmix64(seed ^ (0xD1B54A32D192ED03ull * (unsigned long long)(round + 1)));
    h = splitmix64(h ^ (0x9E3779B97F4A7C15ull * (unsigned long long)(g + 1)));
    h = splitmix64(h ^ (((unsigned long long)j << 32) | (unsigned long long)attempt));
    const int c = (int)(h % (unsigned long long)(n - 1));
    return c + (c >= g ? 1 : 0);
}

__device__ __forceinline__ unsigned load_acquire_sys_u32(const unsigned* addr) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
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
    // ---- all ranks' snapshots of this round are complete (acquire) ----
    if (a.arrive != nullptr) {
        if (tid == 0) while ((int)(load_acquire_sys_u32(a.arrive) - a.expect) < 0) __nanosleep(64);
        __syncthreads();
    }
    if ((int)blockIdx.x >= a.n_sel) return;          // a rank without selected clients still waits (one idle CTA)
    const int c = a.sel_local[blockIdx.x];
    const int g = a.first_client + c;
    const int n = a.n_global;
    if (tid == 0) { mbar_init(&S.row_bar[0], 1); mbar_init(&S.row_bar[1], 1); fence_mbar_init(); }
    // ---- candidate list: self first, then the in-neighbours ----
    if (tid == 0) {
        int m = 0;
        if (a.topo_kind == FM_TOPO_CSR) {
            m = a.nbr_off[g + 1] - a.nbr_off[g] - 1;                  // the table lists self first
        } else if (a.topo_kind == FM_TOPO_FULL) {
            m = n - 1;
        } else if (a.topo_kind == FM_TOPO_RING) {
            const int hops = a.topo_k > 2 ? max(1, a.topo_k / 2) : 1;
            for (int h = 1; h <= hops; ++h) {
                const int cand[2] = {((g - h) % n + n) % n, (g + h) % n};
                for (int s = 0; s < 2; ++s) {
                    bool dup = cand[s] == g;
                    for (int q = 0; q < m; ++q) dup |= S.nbr[q] == cand[s];
                    if (!dup && m < GOSSIP_MAX_K) S.nbr[m++] = cand[s];
                }
            }
        } else {                                                       // FM_TOPO_RANDOM_K
            const int k = min(min(a.topo_k, n - 1), GOSSIP_MAX_K);
            for (int j = 0; j < k; ++j) {
                for (int attempt = 0;; ++attempt) {
                    const int cand = random_neighbour(a.topo_seed, a.round, g, j, attempt, n);
                    bool dup = false;
                    for (int q = 0; q < m; ++q) dup |= S.nbr[q] == cand;
                    if (!dup) { S.nbr[m++] = cand; break; }
                }
            }
        }
        S.nbr[GOSSIP_MAX_K] = m;
    }
    __syncthreads();
    const int nc = S.nbr[GOSSIP_MAX_K];              // neighbours (self excluded)
    auto candidate = [&](int k) -> int {             // k = 0 is the client itself
        if (k == 0) return g;
        if (a.topo_kind == FM_TOPO_CSR) return a.nbr_idx[a.nbr_off[g] + k];
        if (a.topo_kind == FM_TOPO_FULL) return (k - 1) + ((k - 1) >= g ? 1 : 0);
        return S.nbr[k - 1];
    };
    auto fetch = [&](int k) {                        // one elected thread: row k -> raw[k & 1] (unicast bulk copy)
        int r, l;
        owner_of_g(candidate(k), n, a.world, r, l);
        const float* src = a.peer_src[r] + (size_t)a.snap_parity * a.snap_stride + (size_t)l * P_PAD;
        mbar_expect_tx(&S.row_bar[k & 1], P_PAD * 4);
        bulk_g2s(S.raw[k & 1], src, P_PAD * 4, &S.row_bar[k & 1]);
    };
    const float* V = a.x_valid + (size_t)a.valid_off[g] * XP;
    const int nr = a.valid_off[g + 1] - a.valid_off[g];
    for (int i = tid; i < 2 * TILE * XP; i += SNT) (&S.sc.x[0][0])[i] = 0.f;
    for (int p = tid; p < P_PAD; p += SNT) S.theta[p] = 0.f;
    if (tid == 0) fetch(0);
    float ssum = 0.f;
    for (int k = 0; k <= nc; ++k) {
        __syncthreads();                              // everyone is done with raw[(k + 1) & 1] (row k - 1)
        if (tid == 0 && k + 1 <= nc) fetch(k + 1);
        mbar_wait(&S.row_bar[k & 1], (uint32_t)((k >> 1) & 1));
        const float* row = S.raw[k & 1];
        float s = 1.f;
        if (a.mode == 1) {
            load_model_transposed(S.sc, row);
            s = 1.f / model_mse(S, V, nr);
        }
        for (int p = tid; p < P_PAD; p += SNT) S.theta[p] = fmaf(s, row[p], S.theta[p]);
        ssum += s;
    }
    __syncthreads();
    {
        const float inv = 1.f / ssum;
        for (int p = tid; p < P_PAD; p += SNT) S.theta[p] = p < NP ? S.theta[p] * inv : 0.f;
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

// params -> snapshot[parity]; the last block to finish publishes "this rank's snapshot of the round is complete"
// to every rank (release at system scope after a device-wide fence).
__global__ void __launch_bounds__(256) gossip_publish_kernel(const float4* __restrict__ src, float4* __restrict__ dst,
                                                             long long n4, float* const* peer_ctl, int world,
                                                             unsigned* done_counter) {
    for (long long i = (long long)blockIdx.x * blockDim.x + threadIdx.x; i < n4; i += (long long)gridDim.x * blockDim.x)
        dst[i] = src[i];
    if (peer_ctl == nullptr) return;
    __shared__ unsigned last;
    __threadfence();
    __syncthreads();
    if (threadIdx.x == 0) last = atomicAdd(done_counter, 1u);
    __syncthreads();
    if (last == gridDim.x - 1) {
        if (threadIdx.x == 0) *done_counter = 0u;
        __threadfence_system();
        if ((int)threadIdx.x < world) {
            unsigned* arrive = reinterpret_cast<unsigned*>(peer_ctl[threadIdx.x] + 2 * FM_MAX_SEL) + 3;
            asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(arrive) : "memory");
        }
    }
}

}  // namespace fm

extern "C" int fm_gossip(const FmGossipArgs* args, void* stream) {
    using namespace fm;
    if (args->n_sel < 0 || (args->n_sel == 0 && args->arrive == nullptr)) return 0;
    if (args->topo_kind < FM_TOPO_CSR || args->topo_kind > FM_TOPO_RANDOM_K) return (int)cudaErrorInvalidValue;
    if (args->topo_kind == FM_TOPO_CSR && (args->nbr_off == nullptr || args->nbr_idx == nullptr))
        return (int)cudaErrorInvalidValue;
    if ((args->topo_kind == FM_TOPO_RING || args->topo_kind == FM_TOPO_RANDOM_K) &&
        (args->topo_k < 1 || args->topo_k > (args->topo_kind == FM_TOPO_RING ? 2 * GOSSIP_MAX_K : GOSSIP_MAX_K)))
        return (int)cudaErrorInvalidValue;           // the in-kernel list holds 64 neighbours; use a CSR table beyond
    if (args->n_global < 2 && args->topo_kind != FM_TOPO_CSR) return (int)cudaErrorInvalidValue;
    const size_t smem = sizeof(GossipSmem);
    cudaError_t err = cudaFuncSetAttribute(gossip_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    gossip_kernel<<<args->n_sel > 0 ? args->n_sel : 1, SNT, smem, (cudaStream_t)stream>>>(*args);
    return (int)cudaGetLastError();
}

extern "C" int fm_gossip_publish(const float* params, float* snapshot, int64_t n_floats, float* const* peer_ctl,
                                 int32_t world, uint32_t* done_counter, void* stream) {
    if ((n_floats & 3) || world > 32) return (int)cudaErrorInvalidValue;
    const long long n4 = n_floats / 4;
    int grid = (int)((n4 + 255) / 256);
    grid = grid < 1 ? 1 : (grid > 148 * 4 ? 148 * 4 : grid);
    fm::gossip_publish_kernel<<<grid, 256, 0, (cudaStream_t)stream>>>(
        reinterpret_cast<const float4*>(params), reinterpret_cast<float4*>(snapshot), n4, peer_ctl, world, done_counter);
    return (int)cudaGetLastError();
}

extern "C" int32_t fm_random_neighbour(uint64_t seed, int32_t round, int32_t g, int32_t j, int32_t attempt, int32_t n) {
    return fm::random_neighbour(seed, round, g, j, attempt, n);
}
