// Vote scoring as one host call (reference: src/Trainer/client_trainer.py:208-247 for every
// candidate of src/main.py:284-288): normalise the voter's validation matrix, run every selected
// model over it, reduce per 128-row chunk and apply the tie-break factor.  Three launches on the
// caller's stream; the normalised matrix and the per-row errors live in caller-provided scratch.
#include <cuda_runtime.h>
#include <stdio.h>
#include <mutex>
#include "../include/fedmse_b200.h"

extern "C" int fm_vote_scores(const FmVoteArgs* a, void* stream) {
    if (a->n_sel <= 0 || a->rows <= 0) return 0;
    int rc = fm_vote_normalise(a->V, a->rows, a->scratch_norm, stream);
    if (rc) return rc;
    FmScoreArgs s;
    s.slab = a->slab; s.slot_client = a->sel; s.first_client = a->first_client; s.n_local = a->n_local;
    s.X = a->scratch_norm; s.tiles = a->tiles; s.n_tiles = a->n_tiles; s.out_sse = a->scratch_sse; s.out_lat = nullptr;
    rc = fm_score_rows(&s, stream);
    if (rc) return rc;
    return fm_chunk_mean(a->scratch_sse, a->n_sel, a->rows, a->chunk, a->jitter, a->sel, a->first_client,
                         a->n_local, a->voter_pos, a->scores, stream);
}

// ---------------------------------------------------------------------------------------------
// fm_vote_scores_static: the same computation with its scratch inside this library.
//
// Some hosts that want an election score cannot hand us scratch memory because they cannot obtain
// any at that point: a round loop that has promised its operator never to call cudaMalloc after
// start-up (the call serialises the whole device), a process whose GPU memory was claimed up front
// by another framework's pool so that a fresh cudaMalloc would simply fail, or a thin C host that
// links this library and owns no allocator at all.  For them the two scratch buffers of
// fm_vote_scores live in the file-scope device array below.  The array is sized for
// FM_STATIC_VOTE_MAX_ROWS validation rows and FM_STATIC_VOTE_MAX_SEL selected models (768 KiB) and
// every GPU that loads the library keeps it until the process exits.
//
// In host code `g_vote_scratch` is only a handle for the symbol, not an address the GPU can use, so
// the device address is looked up once per device with cudaGetSymbolAddress (with that device
// current) and cached; the launches then receive that address like any caller-owned buffer.
static __device__ __align__(256) float g_vote_scratch[FM_STATIC_VOTE_MAX_ROWS * FM_XP +
                                                      FM_STATIC_VOTE_MAX_SEL * FM_STATIC_VOTE_MAX_ROWS];

namespace {
constexpr int kMaxDevices = 64;
float* g_scratch_addr[kMaxDevices] = {nullptr};
std::mutex g_scratch_mu;
thread_local char g_static_err[256] = "";

// first use on a device: resolve the device address of the scratch array (no allocation involved)
int static_scratch(float** out) {
    int dev = -1;
    cudaError_t e = cudaGetDevice(&dev);
    if (e != cudaSuccess) return (int)e;
    if (dev < 0 || dev >= kMaxDevices) return (int)cudaErrorInvalidDevice;
    std::lock_guard<std::mutex> lock(g_scratch_mu);
    if (g_scratch_addr[dev] == nullptr) {
        float* p = nullptr;
        e = cudaGetSymbolAddress((void**)&p, g_vote_scratch);
        if (e != cudaSuccess) {
            snprintf(g_static_err, sizeof(g_static_err), "cudaGetSymbolAddress(g_vote_scratch) failed on device %d: %s",
                     dev, cudaGetErrorString(e));
            return (int)e;
        }
        g_scratch_addr[dev] = p;
    }
    *out = g_scratch_addr[dev];
    return 0;
}
}  // namespace

extern "C" const char* fm_vote_scores_static_error(void) { return g_static_err; }

extern "C" int fm_vote_scores_static(const FmVoteStaticArgs* a, void* stream) {
    g_static_err[0] = 0;
    if (a->n_sel <= 0 || a->rows <= 0) return 0;
    if (a->rows > FM_STATIC_VOTE_MAX_ROWS || a->n_sel > FM_STATIC_VOTE_MAX_SEL) {
        snprintf(g_static_err, sizeof(g_static_err),
                 "fm_vote_scores_static: %d rows x %d models exceeds the built-in scratch (max %d rows, %d models); "
                 "use fm_vote_scores with caller-provided scratch",
                 a->rows, a->n_sel, FM_STATIC_VOTE_MAX_ROWS, FM_STATIC_VOTE_MAX_SEL);
        return (int)cudaErrorInvalidValue;
    }
    float* base = nullptr;
    int rc = static_scratch(&base);
    if (rc) return rc;
    float* norm = base;                                            // [rows][FM_XP]
    float* sse = base + (size_t)FM_STATIC_VOTE_MAX_ROWS * FM_XP;   // [n_sel][rows]
    // The array still holds whatever the previous call left.  The normalised matrix is rewritten in
    // full below; the per-row errors are zeroed so that a row no tile covers can never contribute a
    // stale value from an earlier, larger call.
    cudaError_t e = cudaMemsetAsync(sse, 0, (size_t)a->n_sel * a->rows * sizeof(float), (cudaStream_t)stream);
    if (e != cudaSuccess) return (int)e;
    FmVoteArgs v;
    v.slab = a->slab; v.sel = a->sel; v.n_sel = a->n_sel; v.first_client = a->first_client; v.n_local = a->n_local;
    v.V = a->V; v.rows = a->rows; v.chunk = a->chunk; v.jitter = a->jitter; v.voter_pos = a->voter_pos;
    v.tiles = a->tiles; v.n_tiles = a->n_tiles; v.scratch_norm = norm; v.scratch_sse = sse; v.scores = a->scores;
    return fm_vote_scores(&v, stream);
}
