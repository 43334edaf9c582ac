/* fedmse_b200 -- C ABI of the sm_100a kernels and the device runtime.
 *
 * Every entry point takes raw device pointers plus the CUDA stream to launch on
 * (as a void*; 0 = default stream), launches asynchronously and returns a
 * cudaError_t value as int (0 = success).  No entry point synchronises the
 * stream.  Unless stated otherwise an entry point allocates nothing: all
 * buffers, including scratch, belong to the caller.  Python binds these with
 * ctypes (fedmse_decentralized_b200/ops/_lib.py); any C/C++ host can link them.
 *
 * Data conventions: every data matrix is fp32 row-major with a row pitch of
 * FM_XP (128) floats, zero padded beyond the 115 features.  A model is a packed
 * fp32 row of FM_P_PAD floats (the 8 state_dict tensors back to back).  Rows of
 * many clients are concatenated and addressed through int32 offset tables
 * off[clients + 1].
 */
#ifndef FEDMSE_B200_H
#define FEDMSE_B200_H

#include <stdint.h>

#ifdef __cplusplus
extern "C" {
#endif

#define FM_D 115
#define FM_H 27
#define FM_Z 7
#define FM_XP 128
#define FM_NP 6764
#define FM_P_PAD 6784
#define FM_MAX_WORLD 16
#define FM_MAX_SEL 4096

/* ---- T1: persistent per-client local training -----------------------------------------
 * One CTA per selected client runs the whole `ClientTrainer.run` call on chip: forward,
 * SAE/AE loss, backward, FedProx term, Adam, per-epoch validation, best-epoch snapshot and
 * early stop (reference: src/Trainer/client_trainer.py:360-419,
 * src/Model/Shrink_Autoencoder.py:152-163).  Weights and Adam moments stay in shared memory
 * for all epochs; batches are streamed with bulk async copies. */
typedef struct FmTrainArgs {
    float* params;            /* [C][P_PAD] in/out */
    float* adam_m;            /* [C][P_PAD] in/out */
    float* adam_v;            /* [C][P_PAD] in/out */
    int32_t* adam_t;          /* [C] in/out optimiser step count */
    const float* anchor;      /* [C][P_PAD] FedProx anchor (read only when mu != 0) */
    float* best;              /* [C][P_PAD] out: weights of the best-validation epoch */
    float* loss_log;          /* [C][log_stride][2] out: (train, valid) loss per epoch */
    int32_t* epochs_run;      /* [C] out */
    int32_t* trained_rounds;  /* [C] in/out, incremented */
    const float* x_train;     /* rows of local clients */
    const int32_t* train_off; /* [C+1] */
    const float* x_valid;     /* rows of all (global) clients */
    const int32_t* valid_off; /* [N_global+1] */
    const int32_t* sel_local; /* [n_sel] local client indices to train */
    int32_t n_sel;
    int32_t first_client;     /* global id of local client 0 */
    int32_t epochs;
    int32_t batch;            /* rows per step, <= 32 */
    int32_t patience;
    int32_t log_stride;
    float lr;
    float shrink_lambda;      /* 0 => plain AE loss */
    float fedprox_mu;         /* 0 => no proximal term */
    unsigned long long* phase_cycles; /* optional [16] debug counters: SM cycles per phase of CTA 0 ([8..10]: warp 0's own work inside three of them), or NULL */
} FmTrainArgs;
int fm_train_clients(const FmTrainArgs* args, void* stream);

/* ---- shape-generic training / scoring (any dim <= FM_XP, hidden, latent; batch <= 4096) ---------
 * The general path behind the specialised kernels: same math and update order as fm_train_clients /
 * fm_score_rows (reference: client_trainer.py:360-419; the models' hidden_neus / latent_dim arguments,
 * Shrink_Autoencoder.py:125-127), parameters laid out as layout.py packs them for (dim, hidden, latent),
 * rows of [C][P_PAD(dim, hidden, latent)].  One CTA per client; weights, moments and activations stay in global
 * memory, so nothing limits the width (the scoring kernel keeps 16 rows of activations in shared memory and
 * serves hidden widths up to ~1,500).  scratch: n_sel * scratch_stride floats with
 * scratch_stride >= fm_train_generic_scratch_floats(dim, hidden, latent, batch). */
typedef struct FmTrainGenericArgs {
    float* params; float* adam_m; float* adam_v; int32_t* adam_t; const float* anchor; float* best;
    float* loss_log; int32_t* epochs_run; int32_t* trained_rounds;
    const float* x_train; const int32_t* train_off; const float* x_valid; const int32_t* valid_off;
    const int32_t* sel_local; int32_t n_sel; int32_t first_client;
    int32_t epochs; int32_t batch; int32_t patience; int32_t log_stride;
    float lr; float shrink_lambda; float fedprox_mu;
    int32_t dim; int32_t hidden; int32_t latent;
    float* scratch; int64_t scratch_stride;
} FmTrainGenericArgs;
int fm_train_generic(const FmTrainGenericArgs* args, void* stream);
int64_t fm_train_generic_scratch_floats(int32_t dim, int32_t hidden, int32_t latent, int32_t batch);

/* tiles[n_tiles][4] = {slot, row0, nrows (<= 16), out0} as for fm_score_rows; out_sse[out0 + r] and / or
 * out_lat[(out0 + r) * lat_pitch + i] are written for slots whose client is local. */
typedef struct FmScoreGenericArgs {
    const float* slab; const int32_t* slot_client; int32_t first_client; int32_t n_local;
    const float* X; const int32_t* tiles; int32_t n_tiles;
    float* out_sse; float* out_lat; int32_t lat_pitch;
    int32_t dim; int32_t hidden; int32_t latent;
} FmScoreGenericArgs;
int fm_score_generic(const FmScoreGenericArgs* args, void* stream);

/* ---- T6: fusion_avg weight function (reference: src/Utils/utils.py:10-24) ---------------------
 * For every slot whose client is local: reconstruct the first n_rows rows of X with the slot's model, take the Gaussian
 * KDE (Scott bandwidth) self log-density of those n_rows points, and write the Jensen-Shannon distance between its
 * exp() and that of the rows themselves to dist_out[slot] (+1e-12, so 1/d is finite).  rec_scratch needs
 * fm_kde_js_scratch_floats(n_slots, n_rows) floats, dens (n_slots + 1) * n_rows floats. */
typedef struct FmKdeJsArgs {
    const float* slab; const int32_t* slot_client; int32_t first_client; int32_t n_local; int32_t n_slots;
    const float* X; int32_t n_rows;
    float* rec_scratch; float* dens; float* dist_out;
} FmKdeJsArgs;
int fm_kde_js(const FmKdeJsArgs* args, void* stream);
int64_t fm_kde_js_scratch_floats(int32_t n_slots, int32_t n_rows);

/* ---- row scoring forward (shared by T2 vote/merge scoring, T3 verification, T4 evaluation) --
 * For every tile {slot, row0, nrows<=64, out0}: run the model of `slot` over rows
 * X[row0 .. row0+nrows) and write per-row sum of squared reconstruction error to
 * out_sse[out0 + r] and/or the latent to out_lat[(out0 + r) * 8 + i].
 * Model of a slot: slot_client ? slab[(slot_client[slot] - first_client)] : slab[slot];
 * slots whose client is not in [first_client, first_client + n_local) are skipped.
 * (reference: the forward passes of client_trainer.py:208-247, :115-130,
 * model_verifier.py:86-99, evaluator.py:56-94). */
typedef struct FmScoreArgs {
    const float* slab;          /* [n_local][P_PAD] (or a single model row with n_local = 1) */
    const int32_t* slot_client; /* [n_slots] global client id per slot, or NULL = identity (local) */
    int32_t first_client;
    int32_t n_local;
    const float* X;             /* data rows, pitch FM_XP */
    const int32_t* tiles;       /* [n_tiles][4] = {slot, row0, nrows, out0} */
    int32_t n_tiles;
    float* out_sse;             /* or NULL */
    float* out_lat;             /* or NULL; 8 floats per row (7 latents + 0) */
} FmScoreArgs;
int fm_score_rows(const FmScoreArgs* args, void* stream);

/* Same contract on the tensor cores (tcgen05.mma kind::tf32, accumulators in TMEM, tiles by TMA):
 * tiles hold up to 128 rows; x_rows = number of rows behind X (rows past it read as zero).
 * precision 1: single TF32 pass (inputs truncated to 10-bit mantissa, fp32 accumulate);
 * precision 3: error-compensated 3xTF32 (hi/lo split of both operands, ~fp32 accuracy);
 * precision 3 reads X as the hi part (the tensor core ignores the low 13 mantissa bits of an fp32
 * operand) and needs x_lo = X - trunc_tf32(X), the residual shard made once by fm_split_residual. */
int fm_score_rows_tc(const FmScoreArgs* args, int64_t x_rows, int32_t precision, const float* x_lo, void* stream);
int fm_split_residual(const float* x, float* x_lo, int64_t rows, void* stream);

/* ---- many models on one shared matrix (MSE-weight scoring on the pooled dev set) -----------
 * Inverts the loop of fm_score_rows_tc: each CTA keeps one 128-row tile of X resident on chip (TF32 head and
 * fp32 remainder in tensor memory as the A operand, the fp32 tile in shared memory for the error) and runs
 * a chunk of the locally owned slots' models over it (operand images pre-staged by a prep kernel and pulled
 * with bulk async copies), so X is read once per unit instead of once per (tile, model).
 * X_lo is ignored since v5 (the split happens in the kernel); the field stays for ABI stability.
 * out_sse is laid out [n_slots][x_rows]; rows of slots owned by other ranks are left untouched.
 * Scratch (caller): images >= owned * FM_TC_IMAGE_BYTES bytes, owned_slots[n_slots] ints, n_owned[1] int.
 * (reference: the full-batch dev-set forward of fed_mse_avg, client_trainer.py:118-124). */
#define FM_TC_IMAGE_BYTES 65536
typedef struct FmMultiScoreArgs {
    const float* slab; const int32_t* slot_client; int32_t n_slots; int32_t first_client; int32_t n_local;
    const float* X; const float* X_lo; int64_t x_rows; int32_t precision;
    unsigned char* images; int32_t* owned_slots; int32_t* n_owned;
    float* out_sse;
    unsigned long long* dbg_cycles; /* optional [32]: CTA 0 cycle breakdown (producer [0..7], epilogue 1 [8..11], epilogue 2 [12..15], clock stamps of the first unit boundaries [16..31]), or NULL */
} FmMultiScoreArgs;
int fm_score_multi_tc(const FmMultiScoreArgs* args, void* stream);

/* ---- MSE-weight scoring, second generation: every layer on tcgen05 with fp16 hi/lo operands ------------
 * Same job as fm_score_multi_tc (reference: fed_mse_avg, client_trainer.py:115-130) but (1) x and the weights as
 * an fp16 head + fp16 remainder (three products in layer 1), the hidden activations as one fp16 (two products in the
 * other GEMMs), fp32 accumulation (kind::f16, twice the TF32 rate; measured ~1e-7 relative on the MSE; operands are
 * clamped to +-65504 first, so data beyond ~1.3e5 in magnitude loses accuracy -- callers with such data use
 * fm_score_multi_tc); (2) the 27->7->27 bottleneck as one fused 27x27 GEMM (W32 = W3 W2, no activation between the
 * two inner layers) and all biases folded into the GEMMs; (3) out_mse[slot] = mean squared error of the slot's model
 * over all x_rows x 115 elements is finished in the kernel (per-thread running sums, added in a fixed order behind a
 * grid-wide barrier): no SSE matrix, no fm_chunk_mean.  Slots owned by other ranks are left untouched.
 * Scratch (caller): images >= owned * FM_TC_IMAGE_BYTES bytes, owned_slots[n_slots], n_owned[1],
 * partial[fm_score_fused_scratch_floats(x_rows, owned)] floats, grid_barrier[2] uint32 zeroed once. */
typedef struct FmFusedScoreArgs {
    const float* slab; const int32_t* slot_client; int32_t n_slots; int32_t first_client; int32_t n_local;
    const float* X; int64_t x_rows;
    unsigned char* images; int32_t* owned_slots; int32_t* n_owned;
    float* partial; float* out_mse; uint32_t* grid_barrier;
    unsigned long long* dbg_cycles;
} FmFusedScoreArgs;
int fm_score_fused_f16(const FmFusedScoreArgs* args, void* stream);
int64_t fm_score_fused_scratch_floats(int64_t x_rows, int32_t n_models);

/* Column-normalise a vote matrix: out = (V - mean) / (unbiased std + 1e-8)
 * (reference: client_trainer.py:221-223).  V and out are [rows][FM_XP]; out is caller scratch. */
int fm_vote_normalise(const float* V, int32_t rows, float* out, void* stream);

/* scores[slot] = jitter[slot] * mean over `chunk`-row blocks of (sum sse / (rows_in_block * D)).
 * sse is laid out [n_slots][rows]; chunk <= 0 means one block.  Slots with skip[slot] != 0 or
 * whose client is not local get 0 (the caller sums over ranks).
 * (reference: client_trainer.py:226-247; with chunk = 0 the full-batch MSE of :123). */
int fm_chunk_mean(const float* sse, int32_t n_slots, int32_t rows, int32_t chunk, const float* jitter,
                  const int32_t* slot_client, int32_t first_client, int32_t n_local, int32_t skip_slot,
                  float* scores, void* stream);

/* ---- T2 vote scoring in one call: fm_vote_normalise + fm_score_rows + fm_chunk_mean ----------
 * (reference: client_trainer.py:208-247 applied to every candidate, main.py:284-288).
 * scores[p] = jitter[p] * chunked MSE of model sel[p] on the normalised V; 0 for the voter's own
 * position and for models owned by other ranks.  tiles must cover, for every position p, rows
 * [0, rows) of the normalised matrix with out0 = p * rows.  scratch_norm holds rows * FM_XP floats,
 * scratch_sse holds n_sel * rows floats; both are caller-allocated and may be reused afterwards. */
typedef struct FmVoteArgs {
    const float* slab; const int32_t* sel; int32_t n_sel; int32_t first_client; int32_t n_local;
    const float* V; int32_t rows; int32_t chunk; const float* jitter; int32_t voter_pos;
    const int32_t* tiles; int32_t n_tiles;
    float* scratch_norm; float* scratch_sse; float* scores;
} FmVoteArgs;
int fm_vote_scores(const FmVoteArgs* args, void* stream);

/* ---- vote scoring without caller scratch --------------------------------------------------
 * No counterpart in the reference (which allocates freely through PyTorch).  Same inputs, outputs
 * and results as fm_vote_scores, for hosts that cannot supply scratch because they cannot allocate
 * device memory at the point of the call: loops that must never hit cudaMalloc's device-wide
 * serialisation after start-up, processes whose GPU memory is already owned by another framework's
 * pool, or plain C hosts without an allocator.  The scratch is a device array built into the
 * library, limited to FM_STATIC_VOTE_MAX_ROWS validation rows and FM_STATIC_VOTE_MAX_SEL selected
 * models (768 KiB, resident on every GPU that loads the library until the process exits).  Larger
 * requests return cudaErrorInvalidValue and leave a message in fm_vote_scores_static_error().
 * The call performs no cudaMalloc / cudaFree and calls no allocating library routine.
 *  - The first call on each device looks up the array's device address; make that first call
 *    outside any stream capture.
 *  - All calls share the one array, which keeps the previous call's contents; the entry point
 *    clears what must start from zero itself.  Two calls whose GPU work can overlap (different
 *    streams, or threads) race on the array: serialise them. */
#define FM_STATIC_VOTE_MAX_ROWS 1024
#define FM_STATIC_VOTE_MAX_SEL 64
typedef struct FmVoteStaticArgs {
    const float* slab; const int32_t* sel; int32_t n_sel; int32_t first_client; int32_t n_local;
    const float* V; int32_t rows; int32_t chunk; const float* jitter; int32_t voter_pos;
    const int32_t* tiles; int32_t n_tiles;
    float* scores;
} FmVoteStaticArgs;
int fm_vote_scores_static(const FmVoteStaticArgs* args, void* stream);
const char* fm_vote_scores_static_error(void);

/* ---- election (reference: client_trainer.py:262-285, main.py:284-288) --------------------
 * agg_pos[0] = argmin over positions p != voter_pos with agg_count[sel[p]] < max_agg of
 * scores[p] (NaN = +inf, ties: first); -1 if none.  agg_pos[1] = the elected global id or -1. */
int fm_elect(const float* scores, const int32_t* sel, int32_t n_sel, int32_t voter_pos,
             const int32_t* agg_count, int32_t max_agg, int32_t* agg_pos, const uint32_t* arrive,
             uint32_t expect, void* stream);

/* ---- cross-GPU scalar exchange without NCCL (multi-GPU election scores / merge MSEs) --------
 * Every rank owns a peer-mapped control block of FM_CTL_FLOATS floats: two boards of FM_MAX_SEL
 * floats (0 = vote scores, 1 = merge MSEs) followed by four uint32 arrival counters (0, 1 = the
 * boards, 2 = "done pulling rows").  fm_publish writes the entries this rank owns into the board
 * `which` of every rank (P2P stores), fences at system scope and bumps arrive[which] on every rank
 * with a release reduction.  A consumer (fm_elect, fm_merge, fm_verify_update) given
 * arrive = &own_ctl.arrive[which] and expect = world * (number of exchanges so far) spins with
 * acquire loads until every rank has arrived and then reads its own board.  Counters only grow. */
#define FM_CTL_FLOATS (2 * FM_MAX_SEL + 16)
int fm_publish(const float* local, const int32_t* slot_client, int32_t first_client, int32_t n_local,
               int32_t n, float* const* peer_ctl, int32_t world, int32_t which, void* stream);

/* ---- T2 merge: weights + weighted parameter reduction with in-kernel peer loads ----------
 * theta = sum_k w_k * model(sel[k]), where model(g) is read from the owner rank's slab through
 * peer_slabs[owner] (P2P over NVLink for remote owners).  Weights: uniform (mode 0) or
 * normalised 1/mse (mode 1), with the reference's aliasing quirk when compat != 0
 * (client_trainer.py:107-130).  Also bumps agg_count[aggregator] (the quota of
 * client_trainer.py:329).  The aggregator's own parameters are replaced by fm_verify_update, after
 * every rank has finished reading the selected rows.  No-op when agg_pos[0] < 0. */
typedef struct FmMergeArgs {
    const float* const* peer_slabs; /* device array [world] of slab base pointers (own rank included) */
    int32_t world;
    int32_t rank;
    int32_t n_global;
    const int32_t* sel;             /* [n_sel] global ids */
    int32_t n_sel;
    const float* mse;               /* [n_sel] or NULL for mode 0 */
    int32_t mode;                   /* 0 avg/fedprox, 1 mse_avg */
    int32_t compat;
    const int32_t* agg_pos;         /* device int[2] from fm_elect */
    int32_t* agg_count;             /* [n_global] */
    float* theta;                   /* [P_PAD] out */
    float* weights_out;             /* [n_sel] out (diagnostics), or NULL */
    const uint32_t* arrive;         /* multi-GPU: wait for arrive[0] >= expect before reading mse; NULL = no wait */
    uint32_t expect;
    float* const* done_ctl;         /* multi-GPU: peers' control blocks; after the reduction every rank bumps */
    uint32_t* done_counter;         /*   arrive[2] on all of them (done_counter: zeroed device scratch); NULL = off */
} FmMergeArgs;
int fm_merge(const FmMergeArgs* args, void* stream);

/* ---- T2 merge through NVLS (multicast reduce inside the NVSwitch) ------------------------------
 * Same result as fm_merge on hosts whose fabric exposes multicast: every rank sums w_k * model(sel[k]) over the
 * selected rows it OWNS into `partial_local` (a symmetric buffer, local reads only), then rank r reduces slice r of
 * all partials with multimem.ld_reduce through `mc_partial` and writes it to every GPU with multimem.st through
 * `mc_theta`; theta (the local view of that buffer) is complete once arrive[FM_ARRIVE_DONE] >= world * seq on
 * this rank (fm_flag_wait orders a stream behind it; fm_verify_update waits for it anyway).
 * Counters: arrive[FM_ARRIVE_PARTIAL] and arrive[FM_ARRIVE_DONE] of the control block of fm_publish are bumped
 * once per call on every rank, also when nothing was elected.  grid_counter: two zeroed uint32 of scratch. */
#define FM_ARRIVE_DONE 2
#define FM_ARRIVE_PARTIAL 4
typedef struct FmMergeNvlsArgs {
    const float* slab;              /* this rank's parameter slab [n_local][P_PAD] */
    int32_t world; int32_t rank; int32_t n_global;
    const int32_t* sel; int32_t n_sel;
    const float* mse; int32_t mode; int32_t compat;
    const int32_t* agg_pos; int32_t* agg_count;
    float* weights_out;
    const uint32_t* arrive_mse; uint32_t expect_mse;          /* NULL for mode 0 */
    float* partial_local;           /* [P_PAD] local view of the symmetric partial buffer */
    const float* mc_partial;        /* multicast view of the same buffer */
    float* mc_theta;                /* multicast view of the symmetric theta buffer */
    float* const* peer_ctl;         /* [world] control blocks */
    const uint32_t* arrive_partial; uint32_t expect_partial;  /* this rank's arrive[FM_ARRIVE_PARTIAL] */
    uint32_t* grid_counter;
} FmMergeNvlsArgs;
int fm_merge_nvls(const FmMergeNvlsArgs* args, void* stream);
int fm_flag_wait(const uint32_t* arrive, uint32_t expect, void* stream);
/* wide models: in-place all-reduce of a symmetric buffer through its multicast view `mc` (rank r reduces and
 * re-broadcasts slice r); waits for `arrive` >= expect first, bumps arrive[which] on every rank when done. */
int fm_nvls_allreduce_slice(float* mc, int32_t world, int32_t rank, int64_t n_floats, const uint32_t* arrive,
                            uint32_t expect, float* const* peer_ctl, int32_t which, uint32_t* grid_counter, void* stream);

/* ---- T3 verification gate + model update (reference: model_verifier.py:29-99,
 * client_trainer.py:174-206).  One CTA per local client c != aggregator:
 *   perf = 1 / (1 + mean(sse rows of c) / D); first time: accept;
 *   else accept iff sum_k ||hist_k - theta_k||_2 <= thr_change and perf - hist_perf >= -thr_perf;
 *   history <- theta (always); on accept params <- theta, anchor <- theta, rejected = 0;
 *   else rejected++.
 * sse rows of client c are sse[sse_off[c] .. sse_off[c+1]) (sse_off may describe one shared
 * segment for all clients). */
typedef struct FmVerifyArgs {
    float* params; float* anchor; float* hist; float* hist_perf; int32_t* hist_valid; int32_t* rejected;
    int32_t n_local; int32_t first_client;
    const float* theta; const int32_t* agg_pos;
    const float* sse; const int32_t* sse_off; int32_t shared_segment; /* 1: every client uses segment 0 */
    float thr_change; float thr_perf;
    float* out_change; float* out_dperf; /* [n_local] diagnostics or NULL */
    const uint32_t* arrive; uint32_t expect; /* multi-GPU: wait until all ranks are done pulling rows; NULL = no wait */
} FmVerifyArgs;
int fm_verify_update(const FmVerifyArgs* args, void* stream);

/* ---- T4 anomaly score -> ROC-AUC (reference: evaluator.py:50-119, Centroid.py:15-35) ------
 * One CTA per local client.  mode 0 (autoencoder): score = sse / D.  mode 1 (hybrid):
 * score = || (z - mean(z_train)) / std(z_train) ||_2 with population std, 0 -> 1.
 * Non-finite scores are mapped like numpy.nan_to_num.  AUC is the tie-aware Mann-Whitney
 * statistic computed exactly in integers after a bitonic sort: in shared memory up to 16,384 test rows
 * per client; above that in `keys_scratch` (n_local * fm_auc_scratch_keys(max_rows) 8-byte words of
 * global memory, required then -- cudaErrorInvalidValue without it). */
typedef struct FmAucArgs {
    int32_t mode; int32_t n_local;
    const float* test_sse; const float* test_lat; const float* labels; const int32_t* test_off;
    const float* train_lat; const int32_t* train_off;
    float* scores_out;  /* [test rows] or NULL */
    double* auc;        /* [n_local] */
    int32_t max_rows;   /* max test rows of any client (sizes shared memory) */
    unsigned long long* keys_scratch;   /* NULL unless fm_auc_scratch_keys(max_rows) > 0 */
} FmAucArgs;
int fm_auc(const FmAucArgs* args, void* stream);
int64_t fm_auc_scratch_keys(int32_t max_rows);

/* ---- gossip merge (per-client T2; reference API client_trainer.py:141-172) ---------------
 * One CTA per selected local client i: candidates = {i} + in-neighbours of i under the topology;
 * weight_j = 1/MSE(model_j on client i's validation rows) (mode 1) or 1 (mode 0), normalised;
 * theta_i = sum w_j model_j, every model fetched ONCE into shared memory (unicast bulk copies through
 * `peer_src`, i.e. NVLink for remote owners) and used for both the score and the reduction; then the
 * verification gate of fm_verify_update on client i's own validation rows.  Reads snapshot buffer
 * `snap_parity` of the `peer_src` slabs, writes `params` (distinct buffers: synchronous gossip).
 * Topology: FM_TOPO_FULL / FM_TOPO_RING (topo_k/2 hops each side) / FM_TOPO_RANDOM_K (topo_k distinct
 * draws of fm_random_neighbour(seed, round, i, slot, attempt, n)) are generated in the kernel;
 * FM_TOPO_CSR reads nbr_off/nbr_idx (global ids, self listed first).
 * Multi-GPU: `arrive` = this rank's arrive[3] counter, complete when >= expect (fm_gossip_publish of
 * every rank); a rank without selected clients still launches one waiting CTA (n_sel = 0). */
enum { FM_TOPO_CSR = 0, FM_TOPO_FULL = 1, FM_TOPO_RING = 2, FM_TOPO_RANDOM_K = 3 };
typedef struct FmGossipArgs {
    const float* const* peer_src;   /* [world] snapshot slabs to read */
    int32_t world; int32_t rank; int32_t n_global;
    float* params; float* anchor; float* hist; float* hist_perf; int32_t* hist_valid; int32_t* rejected;
    int32_t n_local; int32_t first_client;
    const int32_t* sel_local; int32_t n_sel;
    const int32_t* nbr_off; const int32_t* nbr_idx;  /* FM_TOPO_CSR only: CSR over global ids, indexed by global id */
    const float* x_valid; const int32_t* valid_off;
    int32_t mode; float thr_change; float thr_perf;
    int32_t topo_kind; int32_t topo_k; int32_t round; uint64_t topo_seed;
    int32_t snap_parity; int64_t snap_stride;        /* floats between the two snapshot buffers of a slab */
    const uint32_t* arrive; uint32_t expect;         /* NULL on one GPU */
} FmGossipArgs;
int fm_gossip(const FmGossipArgs* args, void* stream);
/* params[n_floats] -> snapshot; then (peer_ctl != NULL) signal arrive[3] of every rank's control block once the
 * whole copy is visible.  done_counter: one zeroed uint32 of scratch owned by the caller. */
int fm_gossip_publish(const float* params, float* snapshot, int64_t n_floats, float* const* peer_ctl,
                      int32_t world, uint32_t* done_counter, void* stream);
/* the random-k draw, for hosts that need the same graph (parallel/topology.py restates it in python) */
int32_t fm_random_neighbour(uint64_t seed, int32_t round, int32_t g, int32_t j, int32_t attempt, int32_t n);

/* ---- wide-model weighted reduction across GPUs by in-kernel peer loads (no NCCL) ------------
 * (BASELINE.json config 5; the reference's merge for a model of any size: client_trainer.py:107-130.)
 * fm_weighted_rows:  out = sum_k w[k] * rows[k], rows = device table of row pointers (local or peer-mapped).
 * fm_slice_reduce:   reduced[slice(rank)] = sum_q partials[q][slice(rank)]      (reduce-scatter by pulling)
 * fm_slice_gather:   out[slice(q)] = reduced[q][slice(q)] for all q             (all-gather by pulling)
 * Optional cross-rank ordering: wait until *arrive >= expect before reading (NULL = no wait); when
 * peer_ctl != NULL the last block bumps arrival counter `which` (index into the uint32 counters that
 * follow the two boards of a control block, see fm_publish) on every rank after its writes are fenced.
 * n_floats must be a multiple of 4 and all buffers 16-B aligned. */
int fm_weighted_rows(const float* const* rows, const float* w, int32_t n, float* out, int64_t n_floats,
                     const uint32_t* arrive, uint32_t expect, float* const* peer_ctl, int32_t world,
                     int32_t which, uint32_t* grid_counter, void* stream);
int fm_slice_reduce(const float* const* partials, int32_t world, int32_t rank, float* reduced, int64_t n_floats,
                    const uint32_t* arrive, uint32_t expect, float* const* peer_ctl, int32_t which,
                    uint32_t* grid_counter, void* stream);
int fm_slice_gather(const float* const* reduced, int32_t world, float* out, int64_t n_floats,
                    const uint32_t* arrive, uint32_t expect, void* stream);

/* ---- runtime: peer-mapped slabs (cudaMalloc + CUDA IPC), used for the parameter slab ------ */
int fm_slab_alloc(uint64_t bytes, void** ptr);                 /* cudaMalloc, zeroed */
int fm_slab_free(void* ptr);
int fm_ipc_get_handle(void* ptr, uint8_t handle[64]);
int fm_ipc_open_handle(const uint8_t handle[64], void** ptr);
int fm_ipc_close_handle(void* ptr);
int fm_enable_peer_access(int32_t peer_device);
const char* fm_error_string(int code);
int fm_device_info(int32_t* sm_count, int32_t* max_smem_optin, int32_t* cc_major, int32_t* cc_minor);

#ifdef __cplusplus
}
#endif
#endif /* FEDMSE_B200_H */
