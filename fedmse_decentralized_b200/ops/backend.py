"""Backend interface of the round engine, and the pure-torch oracle backend.

The round engine (``trainer/engine.py``) is written against seven device-level
operations.  ``CudaBackend`` (``ops/cuda_backend.py``) maps each to a hand-written
sm_100a kernel; ``OracleBackend`` below maps each to the fp32 torch restatement in
``ops/oracle.py`` and runs anywhere (CPU, gloo multi-process).  Both operate on
the same ``FederationState`` / ``ShardStore`` tensors, which is what lets the GPU
tests compare them op by op.

Index conventions: ``ctl.sel`` holds *global* client ids in selection order; a
rank owns the contiguous block ``[first, first + n_local)``.  Per-selection
vectors (scores) are full length ``n_sel``; a rank fills only the positions it
owns and the engine sums them across ranks.
"""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import TYPE_CHECKING, Dict, List, Optional, Sequence, Tuple

import torch

if TYPE_CHECKING:  # avoid the ops <-> trainer import cycle
    from ..data.shards import ShardStore
    from ..trainer.state import FederationState

from . import oracle
from .oracle import TrainHP


@dataclass
class RoundControl:
    """Host-built description of one round (uploaded in one small copy on the CUDA path)."""
    round: int
    sel: List[int]                    # global ids, selection order
    voter_pos: int                    # position in ``sel`` of the effective voter, -1 = nobody can vote
    jitter: List[float]               # tie-break factor per selection position (voter's slot unused)
    vote_client: int                  # global id whose validation matrix is the vote data
    verify_client: int                # compat Q5: global id whose validation matrix verifies everyone; -1 = own


@dataclass
class MergeSpec:
    update_type: str = "avg"
    compat_alias: bool = False        # quirk Q6
    vote_chunk: int = 128
    max_aggregations: int = 3
    verification_threshold: float = 3.0
    performance_threshold: float = 0.002
    verification_method: str = "val"


class OracleBackend:
    name = "oracle"

    def __init__(self, layout, device="cpu"):
        self.layout = layout
        self.device = torch.device(device)

    # ------------------------------------------------------------------ helpers
    @staticmethod
    def _owned(sh: ShardStore, g: int) -> bool:
        return sh.first_client <= g < sh.first_client + sh.n_clients

    # ------------------------------------------------------------------ T1
    def train(self, st: FederationState, sh: ShardStore, sel_local: Sequence[int], hp: TrainHP) -> None:
        st.epochs_run.zero_()
        for c in sel_local:
            t, best, log = oracle.train_client(
                st.params[c], st.adam_m[c], st.adam_v[c], int(st.adam_t[c]), st.anchor[c],
                sh.train.rows(c), sh.valid.rows(sh.first_client + c), self.layout, hp)
            st.adam_t[c] = t
            st.best[c].copy_(best)
            st.loss_log[c].zero_()
            for e, (a, b) in enumerate(log):
                st.loss_log[c, e, 0] = a
                st.loss_log[c, e, 1] = b
            st.epochs_run[c] = len(log)
            st.trained_rounds[c] += 1

    # ------------------------------------------------------------------ T2: vote
    def vote_scores(self, st: FederationState, sh: ShardStore, ctl: RoundControl, spec: MergeSpec
                    ) -> torch.Tensor:
        out = torch.zeros(len(ctl.sel), dtype=torch.float32, device=st.device)
        Vn = oracle.vote_normalise(sh.valid.rows(ctl.vote_client), self.layout.dim)
        for pos, g in enumerate(ctl.sel):
            if pos == ctl.voter_pos or not self._owned(sh, g):
                continue
            s = oracle.chunked_mse(st.params[g - sh.first_client], Vn, self.layout, spec.vote_chunk)
            out[pos] = s * ctl.jitter[pos]
        return out

    def elect(self, st: FederationState, scores: torch.Tensor, ctl: RoundControl, spec: MergeSpec
              ) -> torch.Tensor:
        """argmin over candidates with remaining quota; returns int32[2] = (position, global id) or -1."""
        sel = torch.tensor(ctl.sel, dtype=torch.long, device=scores.device)
        ok = st.agg_count[sel] < spec.max_aggregations
        if ctl.voter_pos >= 0:
            ok[ctl.voter_pos] = False
        else:
            ok[:] = False
        s = torch.where(ok, torch.nan_to_num(scores, nan=float("inf")),
                        torch.full_like(scores, float("inf")))
        # stable: python's sort keeps the first of equal scores
        pos = int(torch.argmin(s)) if bool(ok.any()) else -1
        if pos >= 0 and not bool(ok[pos]):
            pos = int(torch.nonzero(ok)[0])
        return torch.tensor([pos, ctl.sel[pos] if pos >= 0 else -1], dtype=torch.int32, device=scores.device)

    # ------------------------------------------------------------------ T2: merge
    def merge_scores(self, st: FederationState, sh: ShardStore, ctl: RoundControl) -> torch.Tensor:
        out = torch.zeros(len(ctl.sel), dtype=torch.float32, device=st.device)
        for pos, g in enumerate(ctl.sel):
            if self._owned(sh, g):
                out[pos] = oracle.chunked_mse(st.params[g - sh.first_client], sh.dev, self.layout, 0)
        return out

    def fusion_scores(self, st: FederationState, sh: ShardStore, ctl: RoundControl, max_rows: int
                      ) -> torch.Tensor:
        """``fusion_avg``: JS distance between the KDE densities of the dev rows and of each selected model's
        reconstruction of them (``utils.similarity.fusion_distances``); owned positions only, like the MSE."""
        from ..utils.similarity import fusion_distances
        out = torch.zeros(len(ctl.sel), dtype=torch.float32, device=st.device)
        own = [(pos, g - sh.first_client) for pos, g in enumerate(ctl.sel) if self._owned(sh, g)]
        if own:
            d = fusion_distances(st.params[[c for _, c in own]], sh.dev, self.layout, max_rows)
            out[[pos for pos, _ in own]] = d
        return out

    def merge(self, st: FederationState, sh: ShardStore, ctl: RoundControl, spec: MergeSpec,
              mse: Optional[torch.Tensor], agg_pos: torch.Tensor, rows: Optional[torch.Tensor] = None
              ) -> torch.Tensor:
        """Weighted parameter reduction. ``rows`` = the selected models ``[n_sel, PP]`` when some
        live on other ranks (gathered by the engine); ``None`` = all local."""
        a = int(agg_pos[0])
        theta = torch.zeros(self.layout.padded, dtype=torch.float32, device=st.device)
        if a < 0:
            return theta
        n = len(ctl.sel)
        kind = "mse_avg" if spec.update_type == "fusion_avg" else spec.update_type   # weights = (1/d_j) normalised
        w = oracle.merge_weights(kind, mse, n, a, spec.compat_alias, st.device)
        if rows is None:
            rows = st.params[[g - sh.first_client for g in ctl.sel]]
        theta = oracle.weighted_reduce(rows, range(n), w)
        g = ctl.sel[a]
        st.agg_count[g] += 1
        if self._owned(sh, g):
            st.params[g - sh.first_client].copy_(theta)
        return theta

    # ------------------------------------------------------------------ T3
    def verify_update(self, st: FederationState, sh: ShardStore, ctl: RoundControl, spec: MergeSpec,
                      theta: torch.Tensor, agg_pos: torch.Tensor) -> None:
        a = int(agg_pos[0])
        if a < 0:
            return
        agg_g = ctl.sel[a]
        shared_perf = None
        if spec.verification_method == "dev":
            shared_perf = oracle.performance(theta, sh.dev, self.layout)
        elif ctl.verify_client >= 0:
            shared_perf = oracle.performance(theta, sh.valid.rows(ctl.verify_client), self.layout)
        for c in range(st.n_clients):
            g = sh.first_client + c
            if g == agg_g:
                continue
            perf = shared_perf if shared_perf is not None else \
                oracle.performance(theta, sh.valid.rows(g), self.layout)
            if int(st.hist_valid[c]) == 0:
                ok = True
            else:
                change = oracle.param_change(st.hist[c], theta, self.layout)
                delta = perf - st.hist_perf[c]
                ok = bool(change <= spec.verification_threshold) and \
                    bool(delta >= -spec.performance_threshold)
            st.hist[c].copy_(theta)
            st.hist_perf[c] = perf
            st.hist_valid[c] = 1
            if ok:
                st.params[c].copy_(theta)
                st.anchor[c].copy_(theta)
                st.rejected[c] = 0
            else:
                st.rejected[c] += 1

    # ------------------------------------------------------------------ T4
    def evaluate(self, st: FederationState, sh: ShardStore, model_type: str,
                 want_latents: bool = False, metric: str = "AUC"):
        aucs = torch.zeros(st.n_clients, dtype=torch.float32, device=st.device)
        aucs64: List[float] = []
        latents = [] if want_latents else None
        for c in range(st.n_clients):
            score, z = oracle.anomaly_scores(st.params[c], sh.test.rows(c), sh.train.rows(c),
                                             self.layout, model_type)
            if metric == "classification":
                val = oracle.classification_metrics(torch.nan_to_num(score), sh.test.labels(c))[0]
            else:
                val = oracle.auc_from_scores(score, sh.test.labels(c))
            aucs64.append(val)
            aucs[c] = val
            if want_latents:
                latents.append(z.detach().cpu().numpy().astype("float32"))
        return aucs, aucs64, latents

    # ------------------------------------------------------------------ gossip (T2, per-client form)
    def gossip(self, st: FederationState, sh: ShardStore, ctl: RoundControl, spec: MergeSpec,
               neigh: List[List[int]], rows: Optional[torch.Tensor] = None) -> None:
        """Every selected client merges {itself} + its in-neighbours, weighted by 1/MSE of each
        model on the client's *own* validation matrix (uniform for avg/fedprox), then passes the
        result through its verification gate.  Synchronous (Jacobi): all merges read the
        pre-gossip parameters.  ``rows`` = all global models when some are remote."""
        src = (rows if rows is not None else st.params).clone()
        base = 0 if rows is not None else sh.first_client
        for g in ctl.sel:
            if not self._owned(sh, g):
                continue
            c = g - sh.first_client
            cand = [g] + [j for j in neigh[g] if j != g]
            V = sh.valid.rows(g)
            if spec.update_type == "fusion_avg":
                from ..utils.similarity import fusion_distances
                dist_ = fusion_distances(src[[j - base for j in cand]], V, self.layout)
                w = oracle.merge_weights("mse_avg", dist_, len(cand), device=st.device)
            elif spec.update_type == "mse_avg":
                mse = torch.stack([oracle.chunked_mse(src[j - base], V, self.layout, 0) for j in cand])
                w = oracle.merge_weights("mse_avg", mse, len(cand), device=st.device)
            else:
                w = oracle.merge_weights("avg", None, len(cand), device=st.device)
            theta = oracle.weighted_reduce(src, [j - base for j in cand], w)
            perf = oracle.performance(theta, V, self.layout)
            if int(st.hist_valid[c]) == 0:
                ok = True
            else:
                change = oracle.param_change(st.hist[c], theta, self.layout)
                ok = bool(change <= spec.verification_threshold) and \
                    bool(perf - st.hist_perf[c] >= -spec.performance_threshold)
            st.hist[c].copy_(theta)
            st.hist_perf[c] = perf
            st.hist_valid[c] = 1
            if ok:
                st.params[c].copy_(theta)
                st.anchor[c].copy_(theta)
                st.rejected[c] = 0
            else:
                st.rejected[c] += 1

    def synchronize(self) -> None:
        pass
