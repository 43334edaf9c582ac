// T2 merge / election / T3 verification gate.
//
// Reference: election src/Trainer/client_trainer.py:262-285 (+ src/main.py:284-288); merge rules
// :107-134; verification src/Trainer/model_verifier.py:29-99 and client_trainer.py:174-206.
//
// The merge kernel turns MSE scores into normalised weights and performs the weighted parameter
// reduction in one launch.  Every selected model is read straight from its owner's parameter slab
// through a table of peer-mapped base pointers: for models owned by another GPU these are plain
// ld.global on NVLink-mapped addresses (P2P over NVSwitch), so no all-gather, staging buffer or
// separate elementwise kernel exists on this path.  Each rank computes the full reduction for itself
// (an all-reduce by pull), which also *is* the broadcast of the aggregated model.
#include "common.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

__device__ __forceinline__ void owner_of(int g, int n_global, int world, int& rank, int& local) {
    const int base = n_global / world, rem = n_global - base * world;
    const int cut = rem * (base + 1);
    if (g < cut) { rank = g / (base + 1); local = g - rank * (base + 1); }
    else { const int q = (g - cut) / base; rank = rem + q; local = (g - cut) - q * base; }
}

// ---- cross-GPU scalar exchange through peer-mapped control blocks (no NCCL on the merge path) ----
// Every rank owns a control block {float board[2][FM_MAX_SEL]; unsigned arrive[4]} in peer-mapped memory.
// A producer writes the entries it owns into the board of EVERY rank (plain stores over NVLink), fences at
// system scope, and bumps arrive[which] on every rank with a release reduction; a consumer kernel spins with
// acquire loads on its own arrive[which] until all ranks of this exchange have arrived, then reads its board.
__device__ __forceinline__ void signal_release_add(unsigned* addr) {
    asm volatile("red.release.sys.global.add.u32 [%0], 1;" ::"l"(addr) : "memory");
}
__device__ __forceinline__ unsigned load_acquire_sys(const unsigned* addr) {
    unsigned v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(addr) : "memory");
    return v;
}
// block-wide: returns once arrive >= expect (thread 0 polls; everyone else waits at the barrier)
__device__ __forceinline__ void wait_arrivals(const unsigned* arrive, unsigned expect) {
    if (arrive == nullptr) return;
    if (threadIdx.x == 0) {
        while ((int)(load_acquire_sys(arrive) - expect) < 0) __nanosleep(64);
    }
    __syncthreads();
}

constexpr int CTL_BOARD_FLOATS = 2 * FM_MAX_SEL;      // two boards (vote, mse)
__device__ __forceinline__ float* ctl_board(float* ctl, int which) { return ctl + which * FM_MAX_SEL; }
__device__ __forceinline__ unsigned* ctl_arrive(float* ctl, int which) {
    return reinterpret_cast<unsigned*>(ctl + CTL_BOARD_FLOATS) + which;
}

__global__ void __launch_bounds__(256) publish_kernel(const float* __restrict__ local, const int* __restrict__ slot_client,
                                                      int first_client, int n_local, int n, float* const* peer_ctl,
                                                      int world, int which) {
    for (int slot = threadIdx.x; slot < n; slot += blockDim.x) {
        const int c = (slot_client ? slot_client[slot] : first_client + slot) - first_client;
        if (c < 0 || c >= n_local) continue;          // not mine: its owner publishes it
        const float v = local[slot];
        for (int r = 0; r < world; ++r) ctl_board(peer_ctl[r], which)[slot] = v;
    }
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x < world) signal_release_add(ctl_arrive(peer_ctl[threadIdx.x], which));
}

__global__ void elect_kernel(const float* __restrict__ scores, const int* __restrict__ sel, int n_sel, int voter_pos,
                             const int* __restrict__ agg_count, int max_agg, int* __restrict__ agg_pos,
                             const unsigned* arrive, unsigned expect) {
    // single warp: lexicographic (score, position) argmin over eligible candidates
    const int lane = threadIdx.x;
    wait_arrivals(arrive, expect);        // multi-GPU: scores is this rank's board, complete once all ranks arrived
    float best = __int_as_float(0x7f800000);
    int bpos = 0x7fffffff;
    if (voter_pos >= 0) {
        for (int p = lane; p < n_sel; p += 32) {
            if (p == voter_pos || agg_count[sel[p]] >= max_agg) continue;
            float s = scores[p];
            if (s != s) s = __int_as_float(0x7f800000);
            if (s < best || (s == best && p < bpos)) { best = s; bpos = p; }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        const float ob = __shfl_xor_sync(0xffffffffu, best, o);
        const int op = __shfl_xor_sync(0xffffffffu, bpos, o);
        if (ob < best || (ob == best && op < bpos)) { best = ob; bpos = op; }
    }
    if (lane == 0) {
        const bool ok = bpos != 0x7fffffff;
        agg_pos[0] = ok ? bpos : -1;
        agg_pos[1] = ok ? sel[bpos] : -1;
    }
}

constexpr int MNT = 256;

__global__ void __launch_bounds__(MNT) merge_kernel(const FmMergeArgs a) {
    extern __shared__ __align__(16) unsigned char msm[];
    const float** srcs = reinterpret_cast<const float**>(msm);                       // [n_sel] source rows
    float* wts = reinterpret_cast<float*>(msm + (size_t)a.n_sel * sizeof(float*));   // [n_sel]
    __shared__ double red[34];
    const int ap = a.agg_pos[0];
    const int tid = threadIdx.x, n = a.n_sel;
    if (ap < 0) {                         // nothing to merge; keep the arrival counters in step with the host
        if (a.done_ctl != nullptr && blockIdx.x == 0 && tid < a.world) signal_release_add(ctl_arrive(a.done_ctl[tid], 2));
        return;
    }
    wait_arrivals(a.arrive, a.expect);    // multi-GPU: a.mse is this rank's board
    for (int k = tid; k < n; k += MNT) {
        int r, l;
        owner_of(a.sel[k], a.n_global, a.world, r, l);
        srcs[k] = a.peer_slabs[r] + (size_t)l * P_PAD;
    }
    // ---- weights (every CTA recomputes them: n_sel is small) ----
    if (a.mode == 0) {
        for (int k = tid; k < n; k += MNT) wts[k] = 1.f / (float)n;
        __syncthreads();
    } else {
        double part = 0.0;
        for (int k = tid; k < n; k += MNT) {
            int src = k;
            if (a.compat && n > 1 && k == ap && ap > 0) src = ap - 1;   // scored while holding model ap-1
            const float s = 1.f / a.mse[src];
            wts[k] = s;
            part += (double)s;
        }
        const double tot = block_sum<double>(part, red);
        for (int k = tid; k < n; k += MNT) wts[k] = (float)((double)wts[k] / tot);
        __syncthreads();
        if (a.compat && n > 1 && tid == 0) {                            // aggregator's slot aliases the last-loaded peer
            const int last = (ap != n - 1) ? n - 1 : n - 2;
            const float moved = wts[ap];
            wts[ap] = 0.f;
            wts[last] += moved;
        }
        __syncthreads();
    }
    if (a.weights_out && blockIdx.x == 0)
        for (int k = tid; k < n; k += MNT) a.weights_out[k] = wts[k];
    // ---- weighted reduction, one parameter per thread, models pulled from their owners ----
    const int p = blockIdx.x * MNT + tid;
    if (p < P_PAD) {
        float acc = 0.f;
        if (p < NP) {
            int k = 0;
            for (; k + 8 <= n; k += 8) {            // 8 independent (possibly peer) loads in flight per thread
                float v[8];
#pragma unroll
                for (int u = 0; u < 8; ++u) v[u] = srcs[k + u][p];
#pragma unroll
                for (int u = 0; u < 8; ++u) acc = fmaf(wts[k + u], v[u], acc);
            }
            for (; k < n; ++k) acc = fmaf(wts[k], srcs[k][p], acc);
        }
        a.theta[p] = acc;
    }
    if (blockIdx.x == 0 && tid == 0) a.agg_count[a.sel[ap]] += 1;
    // tell every rank that this rank has finished pulling rows: owners may overwrite their slabs afterwards
    if (a.done_ctl != nullptr) {
        __shared__ unsigned last;
        __threadfence();
        __syncthreads();
        if (tid == 0) last = atomicAdd(a.done_counter, 1u);
        __syncthreads();
        if (last == gridDim.x - 1) {
            if (tid == 0) *a.done_counter = 0u;
            if (tid < a.world) signal_release_add(ctl_arrive(a.done_ctl[tid], 2));
        }
    }
}

constexpr int VNT = 256;

__device__ __forceinline__ void verify_decide_and_apply(float* params_c, float* anchor_c, float* hist_c, float* hist_perf,
                                                         int* hist_valid, int* rejected, int c, const float* theta,
                                                         float perf, float thr_change, float thr_perf, double* red,
                                                         float* out_change, float* out_dperf) {
    const int tid = threadIdx.x;
    const bool first = hist_valid[c] == 0;
    float change = 0.f, dperf = 0.f;
    bool ok = true;
    if (!first) {
        float part[8] = {0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f, 0.f};
        for (int p = tid; p < NP; p += VNT) {
            const float dl = hist_c[p] - theta[p];
            part[tensor_of(p)] = fmaf(dl, dl, part[tensor_of(p)]);
        }
#pragma unroll
        for (int t = 0; t < 8; ++t) change += sqrtf((float)block_sum<double>((double)part[t], red));
        dperf = perf - hist_perf[c];
        ok = (change <= thr_change) && (dperf >= -thr_perf);
    }
    __syncthreads();
    for (int p = tid; p < P_PAD; p += VNT) {
        const float t = theta[p];
        hist_c[p] = t;
        if (ok) { params_c[p] = t; anchor_c[p] = t; }
    }
    if (tid == 0) {
        hist_perf[c] = perf;
        hist_valid[c] = 1;
        rejected[c] = ok ? 0 : rejected[c] + 1;
        if (out_change) out_change[c] = change;
        if (out_dperf) out_dperf[c] = dperf;
    }
}

__global__ void __launch_bounds__(VNT) verify_kernel(const FmVerifyArgs a) {
    __shared__ double red[34];
    // multi-GPU: every rank is done reading the pre-merge slabs and score boards (waited for even when nothing was
    // elected: merge_kernel signals in that case too, and the wait keeps the boards' write-after-read order)
    wait_arrivals(a.arrive, a.expect);
    if (a.agg_pos[0] < 0) return;
    const int c = blockIdx.x, tid = threadIdx.x;
    const int g = a.first_client + c;
    float* params_c = a.params + (size_t)c * P_PAD;
    if (g == a.agg_pos[1]) {                       // the aggregator just adopts its own result (no gate, anchor kept)
        for (int p = tid; p < P_PAD; p += VNT) params_c[p] = a.theta[p];
        return;
    }
    const int seg = a.shared_segment ? 0 : c;
    const int r0 = a.sse_off[seg], nr = a.sse_off[seg + 1] - r0;
    double part = 0.0;
    for (int r = tid; r < nr; r += VNT) part += (double)a.sse[r0 + r];
    const double tot = block_sum<double>(part, red);
    const float mse = (float)(tot / ((double)nr * (double)D));
    const float perf = 1.f / (1.f + mse);
    verify_decide_and_apply(params_c, a.anchor + (size_t)c * P_PAD, a.hist + (size_t)c * P_PAD, a.hist_perf,
                            a.hist_valid, a.rejected, c, a.theta, perf, a.thr_change, a.thr_perf, red,
                            a.out_change, a.out_dperf);
}

}  // namespace fm

extern "C" int fm_elect(const float* scores, const int32_t* sel, int32_t n_sel, int32_t voter_pos,
                        const int32_t* agg_count, int32_t max_agg, int32_t* agg_pos, const uint32_t* arrive,
                        uint32_t expect, void* stream) {
    fm::elect_kernel<<<1, 32, 0, (cudaStream_t)stream>>>(scores, sel, n_sel, voter_pos, agg_count, max_agg, agg_pos,
                                                         arrive, expect);
    return (int)cudaGetLastError();
}

extern "C" int fm_merge(const FmMergeArgs* args, void* stream) {
    using namespace fm;
    if (args->n_sel <= 0 || args->n_sel > FM_MAX_SEL) return (int)cudaErrorInvalidValue;
    const int grid = (P_PAD + MNT - 1) / MNT;
    merge_kernel<<<grid, MNT, args->n_sel * (sizeof(float) + sizeof(float*)), (cudaStream_t)stream>>>(*args);
    return (int)cudaGetLastError();
}

extern "C" int fm_verify_update(const FmVerifyArgs* args, void* stream) {
    using namespace fm;
    if (args->n_local <= 0) return 0;
    verify_kernel<<<args->n_local, VNT, 0, (cudaStream_t)stream>>>(*args);
    return (int)cudaGetLastError();
}

extern "C" int fm_publish(const float* local, const int32_t* slot_client, int32_t first_client, int32_t n_local,
                          int32_t n, float* const* peer_ctl, int32_t world, int32_t which, void* stream) {
    if (world > 32 || which < 0 || which > 1 || n > FM_MAX_SEL) return (int)cudaErrorInvalidValue;
    fm::publish_kernel<<<1, 256, 0, (cudaStream_t)stream>>>(local, slot_client, first_client, n_local, n, peer_ctl,
                                                            world, which);
    return (int)cudaGetLastError();
}
