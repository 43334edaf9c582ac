// Row-scoring forward (exact fp32 path): model forward over many data rows, emitting the per-row
// reconstruction SSE and/or the latent.  This is the workhorse behind T2 (vote and MSE-weight
// scoring: src/Trainer/client_trainer.py:208-247, :115-130), T3 (verification forward:
// src/Trainer/model_verifier.py:86-99) and T4 (evaluation: src/Evaluator/evaluator.py:56-94).
//
// The reference runs these as Python loops over batches of 12 (evaluation) or 128 rows (vote) per
// client, one launch-set per batch.  Here one persistent launch covers every (model, row tile) of a
// phase: a tile table drives CTAs over 64-row tiles, tiles are streamed by the TMA unit (bulk async
// copy + mbarrier, double buffered), each CTA keeps the current model transposed in shared memory
// and only reloads it when its tile stream moves to another model.
#include "score_core.cuh"
#include "../include/fedmse_b200.h"

namespace fm {

__global__ void __launch_bounds__(SNT, 2) score_rows_kernel(const FmScoreArgs a, int tiles_per_cta) {
    extern __shared__ __align__(128) unsigned char smem_raw[];
    ScoreSmem& S = *reinterpret_cast<ScoreSmem*>(smem_raw);
    const int tid = threadIdx.x, lane = tid & 31, w = tid >> 5;
    if (blockIdx.x * tiles_per_cta >= a.n_tiles) return;
    for (int i = tid; i < 2 * TILE * XP; i += SNT) (&S.x[0][0])[i] = 0.f;
    if (tid == 0) { mbar_init(&S.bar[0], 1); mbar_init(&S.bar[1], 1); fence_mbar_init(); }
    fence_proxy_async();
    __syncthreads();

    auto local_of = [&](int slot) -> int {      // local model row or -1 (not owned)
        const int g = a.slot_client ? a.slot_client[slot] : a.first_client + slot;
        const int c = g - a.first_client;
        return (c >= 0 && c < a.n_local) ? c : -1;
    };
    auto issue = [&](int t, int buf) {           // thread 0
        const int4 tl = reinterpret_cast<const int4*>(a.tiles)[t];
        if (local_of(tl.x) < 0) return false;
        const uint32_t bytes = (uint32_t)tl.z * XP * 4u;
        mbar_expect_tx(&S.bar[buf], bytes);
        bulk_g2s(&S.x[buf][0], a.X + (size_t)tl.y * XP, bytes, &S.bar[buf]);
        return true;
    };
    uint32_t uses0 = 0, uses1 = 0;
    int cur_model = -1;
    // units of `tiles_per_cta` consecutive tiles dealt round-robin to CTAs (balances ranks that own 1/world of the slots)
    for (int t_begin = blockIdx.x * tiles_per_cta; t_begin < a.n_tiles; t_begin += gridDim.x * tiles_per_cta) {
    const int t_end = min(a.n_tiles, t_begin + tiles_per_cta);
    if (tid == 0) issue(t_begin, 0);
    for (int t = t_begin; t < t_end; ++t) {
        const int buf = (t - t_begin) & 1;
        const int4 tl = reinterpret_cast<const int4*>(a.tiles)[t];
        const int c = local_of(tl.x);
        if (tid == 0 && t + 1 < t_end) issue(t + 1, buf ^ 1);
        if (c < 0) continue;                      // model lives on another rank: its owner scores it
        if (c != cur_model) { load_model_transposed(S, a.slab + (size_t)c * P_PAD); cur_model = c; }
        mbar_wait(&S.bar[buf], (buf ? uses1 : uses0) & 1u);
        if (buf) ++uses1; else ++uses0;
        const bool need_sse = a.out_sse != nullptr;
#pragma unroll 1
        for (int grp = 0; grp < 2; ++grp) {
            const int r0 = w * 8 + grp * 4;
            if (r0 >= tl.z) break;
            float sse[4], zl[4];
            score4(S, &S.x[buf][r0 * XP], w, lane, need_sse, sse, zl);
            if (a.out_lat && lane < 4 * Z) {
                const int r = lane / Z, i = lane - r * Z;
                if (r0 + r < tl.z) a.out_lat[(size_t)(tl.w + r0 + r) * 8 + i] = zl[0];
            }
            if (a.out_lat && lane < 4 && r0 + lane < tl.z) a.out_lat[(size_t)(tl.w + r0 + lane) * 8 + 7] = 0.f;
            if (need_sse && lane < 4 && r0 + lane < tl.z)
                a.out_sse[tl.w + r0 + lane] = lane == 0 ? sse[0] : lane == 1 ? sse[1] : lane == 2 ? sse[2] : sse[3];
        }
        __syncthreads();   // everyone is done with x[buf] before it is refilled two tiles later
    }
    }
}

// ---- vote matrix normalisation: block = 32 columns x 8 row lanes -----------------------------
__global__ void vote_normalise_kernel(const float* __restrict__ V, int rows, float* __restrict__ out) {
    __shared__ double red[8][33];
    const int tx = threadIdx.x, ty = threadIdx.y;
    const int col = blockIdx.x * 32 + tx;
    const bool live = col < D;
    double s = 0.0;
    if (live) for (int r = ty; r < rows; r += 8) s += (double)V[(size_t)r * XP + col];
    red[ty][tx] = s;
    __syncthreads();
    double mean = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) mean += red[i][tx];
    mean /= rows;
    __syncthreads();
    double q = 0.0;
    if (live) for (int r = ty; r < rows; r += 8) { const double dl = (double)V[(size_t)r * XP + col] - mean; q += dl * dl; }
    red[ty][tx] = q;
    __syncthreads();
    double var = 0.0;
#pragma unroll
    for (int i = 0; i < 8; ++i) var += red[i][tx];
    const float meanf = (float)mean;
    const float stdf = (float)sqrt(var / (double)(rows - 1)) + 1e-8f;   // unbiased, as torch.std
    for (int r = ty; r < rows; r += 8)
        out[(size_t)r * XP + col] = live ? (V[(size_t)r * XP + col] - meanf) / stdf : 0.f;
}

// ---- chunked mean of per-row SSE: one block per slot ---------------------------------------------
__global__ void __launch_bounds__(256) chunk_mean_kernel(const float* __restrict__ sse, int n_slots, int rows, int chunk,
                                  const float* __restrict__ jitter, const int* __restrict__ slot_client,
                                  int first_client, int n_local, int skip_slot, float* __restrict__ scores) {
    __shared__ double red[34];
    const int slot = blockIdx.x, tid = threadIdx.x;
    const int g = slot_client ? slot_client[slot] : first_client + slot;
    const int c = g - first_client;
    if (slot == skip_slot || c < 0 || c >= n_local) { if (tid == 0) scores[slot] = 0.f; return; }
    if (chunk <= 0 || chunk > rows) chunk = rows;
    const float* s = sse + (size_t)slot * rows;
    float acc = 0.f; int nchunks = 0;
    for (int r0 = 0; r0 < rows; r0 += chunk, ++nchunks) {
        const int n = min(chunk, rows - r0);
        double part = 0.0;
        for (int r = tid; r < n; r += 256) part += (double)s[r0 + r];
        const double tot = block_sum<double>(part, red);
        acc += (float)(tot / ((double)n * (double)D));
    }
    if (tid == 0) scores[slot] = (acc / (float)nchunks) * (jitter ? jitter[slot] : 1.f);
}

}  // namespace fm

extern "C" int fm_score_rows(const FmScoreArgs* args, void* stream) {
    using namespace fm;
    if (args->n_tiles <= 0) return 0;
    static int sms = 0;
    if (!sms) { int dev; cudaGetDevice(&dev); cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev); }
    const size_t smem = sizeof(ScoreSmem);
    cudaError_t err = cudaFuncSetAttribute(score_rows_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
    if (err != cudaSuccess) return (int)err;
    int per = (args->n_tiles + 2 * sms - 1) / (2 * sms);
    if (per > 32) per = 32;
    int grid = (args->n_tiles + per - 1) / per;
    if (grid > 2 * sms) grid = 2 * sms;
    score_rows_kernel<<<grid, SNT, smem, (cudaStream_t)stream>>>(*args, per);
    return (int)cudaGetLastError();
}

extern "C" int fm_vote_normalise(const float* V, int32_t rows, float* out, void* stream) {
    if (rows <= 0) return 0;
    fm::vote_normalise_kernel<<<4, dim3(32, 8), 0, (cudaStream_t)stream>>>(V, rows, out);
    return (int)cudaGetLastError();
}

extern "C" int fm_chunk_mean(const float* sse, int32_t n_slots, int32_t rows, int32_t chunk, const float* jitter,
                             const int32_t* slot_client, int32_t first_client, int32_t n_local, int32_t skip_slot,
                             float* scores, void* stream) {
    if (n_slots <= 0) return 0;
    fm::chunk_mean_kernel<<<n_slots, 256, 0, (cudaStream_t)stream>>>(
        sse, n_slots, rows, chunk, jitter, slot_client, first_client, n_local, skip_slot, scores);
    return (int)cudaGetLastError();
}
