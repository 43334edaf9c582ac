// Vote scoring as one host call (reference: src/Trainer/client_trainer.py:208-247 for every
// candidate of src/main.py:284-288): normalise the voter's validation matrix, run every selected
// model over it, reduce per 128-row chunk and apply the tie-break factor.  Three launches on the
// caller's stream; the normalised matrix and the per-row errors live in caller-provided scratch.
#include <cuda_runtime.h>
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
