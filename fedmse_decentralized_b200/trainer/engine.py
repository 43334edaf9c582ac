"""Round engine: selection -> local training -> election -> merge -> broadcast/verify -> evaluate.

This is the framework's equivalent of the reference's top-level round loop
(``src/main.py:267-365``) and of the peer protocol methods of ``ClientTrainer``
(``src/Trainer/client_trainer.py:249-335``), restructured so that the host
touches the device O(1) times per round instead of O(clients x steps x ops):

* the host decides only what needs no device data: which clients are selected
  (python RNG, ``src/main.py:270-271``), which selected client is the effective
  voter (a function of the aggregation quotas alone) and the tie-break jitter
  draws (``client_trainer.py:243-245``);
* everything else -- local training of all selected clients, vote scoring, the
  argmin election with quota, merge weights, the weighted parameter reduction,
  the verification gate for every client and the evaluation of every client --
  runs on the device through ``ops.backend`` and is read back once per round.

Two aggregation modes: ``elected`` (the reference's algorithm: one peer-elected
aggregator per round, result broadcast to *all* clients and gated by each
client's verifier) and ``gossip`` (every selected client merges its topology
neighbours' models weighted by their MSE on its own validation data -- the
all-to-all API the reference defines but never calls, quirk Q21).
"""
from __future__ import annotations

import logging
import os
import random
from dataclasses import dataclass, field
from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
import torch

from ..config import ExperimentConfig
from ..data.shards import ShardStore
from ..layout import ParamLayout
from ..models import init_packed_models
from ..ops.backend import MergeSpec, OracleBackend, RoundControl
from ..ops.oracle import TrainHP
from ..parallel import Comm, topology
from ..utils import artefacts
from ..utils.timers import PhaseTimer
from .state import FederationState

log = logging.getLogger(__name__)


@dataclass
class RoundResult:
    round: int
    selected: List[int]
    aggregator: int                        # global client id, -1 = no aggregation
    client_metrics: List[float]            # all global clients
    verification: List[Dict]               # [{'client_id','rejected_updates','is_verified'}]
    epochs_run: Dict[int, int]             # global id -> epochs of this round's local training
    samples: int = 0                       # local training rows processed (train+valid) x epochs


class GlobalStop:
    """The reference's global early stop (``src/main.py:55-57``, ``:358-365``).

    "loss" is ``min(client_metrics)``, i.e. the worst client's AUC, and *lower is
    better*; the counters are module globals that are never reset between
    combinations (quirk Q2).  ``reference_compat`` keeps one instance for the whole
    sweep; otherwise the runner makes a fresh one per combination.
    """

    def __init__(self, patience: int = 1):
        self.patience = patience
        self.min_val_loss = float("inf")
        self.worse = 0

    def update(self, client_metrics: Sequence[float]) -> bool:
        cur = min(client_metrics)
        if cur < self.min_val_loss:
            self.min_val_loss, self.worse = cur, 0
            return False
        self.worse += 1
        return self.worse > self.patience


def make_backend(name: str, layout: ParamLayout, device: torch.device, score_precision: str = "fp32"):
    if name == "auto":
        name = "cuda" if device.type == "cuda" else "oracle"
    if name == "oracle":
        return OracleBackend(layout, device)
    if name == "cuda":
        if (layout.dim, layout.hidden, layout.latent) != (115, 27, 7):
            # widened / narrowed autoencoders: shape-generic kernels (the specialised ones are laid out for 115-27-7)
            from ..ops.generic_backend import GenericCudaBackend
            return GenericCudaBackend(layout, device)
        from ..ops.cuda_backend import CudaBackend
        return CudaBackend(layout, device, score_precision)
    raise ValueError(f"unknown backend {name!r}")


class FederationEngine:
    def __init__(self, cfg: ExperimentConfig, shards: ShardStore, model_type: str, update_type: str,
                 run: int = 0, comm: Optional[Comm] = None, backend=None,
                 init_params: Optional[torch.Tensor] = None, names_global: Optional[List[str]] = None):
        self.cfg, self.sh, self.model_type, self.update_type, self.run = cfg, shards, model_type, update_type, run
        self.comm = comm or Comm(device=shards.device)
        self.layout = ParamLayout(cfg.dim_features, cfg.hidden_neus, cfg.latent_dim)
        self.N = shards.n_global
        self.first, self.n_local = shards.first_client, shards.n_clients
        self.names_global = names_global or list(shards.names)
        dev = shards.device
        if init_params is None:
            # same RNG order as the reference: one module per client on the global torch generator
            init_params = init_packed_models(self.N, model_type, self.layout, cfg.shrink_lambda)
        self.backend = backend if backend is not None and not isinstance(backend, str) else \
            make_backend(backend or cfg.backend, self.layout, dev, cfg.score_precision)
        storage = None
        if hasattr(self.backend, "alloc_param_slab"):
            storage = self.backend.alloc_param_slab(self.n_local, self.layout.padded, self.comm)
        self.state = FederationState.create(
            init_params[self.first:self.first + self.n_local], self.layout, cfg.epoch, dev,
            params_storage=storage, n_global=self.N)
        self.hp = TrainHP(
            epochs=cfg.epoch, batch_size=cfg.batch_size, lr=cfg.lr_rate,
            shrink_lambda=float(cfg.shrink_lambda) if model_type == "hybrid" else 0.0,
            fedprox_mu=cfg.fedprox_mu if update_type == "fedprox" else 0.0,
            patience=cfg.local_patience)
        self.spec = MergeSpec(
            update_type=update_type,
            compat_alias=cfg.reference_compat and update_type == "mse_avg",
            vote_chunk=cfg.vote_chunk, max_aggregations=cfg.max_aggregation_threshold,
            verification_threshold=cfg.verification_threshold,
            performance_threshold=cfg.performance_threshold,
            verification_method=cfg.verification_method)
        self.agg_count_host = [0] * self.N
        self.timer = PhaseTimer(enabled=True, cuda=dev.type == "cuda")
        self.latents: Dict[int, Dict[str, Tuple[np.ndarray, np.ndarray]]] = {}
        self.last_metrics: Optional[List[float]] = None
        self.history: List[RoundResult] = []
        self._neigh_cache = None
        self._rand_fifo: List[float] = []
        self._prefetched = None
        self._pre = None                 # (round, control) whose upload + local training are already enqueued
        if hasattr(self.backend, "bind"):
            self.backend.bind(self.state, self.sh, self.comm, self.hp, self.spec, model_type)

    # ------------------------------------------------------------------ control (host RNG)
    def _next_rand(self) -> float:
        """Next ``torch.rand(1).item()`` of the reference's stream (possibly drawn ahead of time)."""
        if self._rand_fifo:
            return self._rand_fifo.pop(0)
        return torch.rand(1).item()

    def prefetch_control(self, round_idx: int) -> None:
        """Do the host work of the NEXT round's control block while the GPU is busy with this one: the
        client selection and the uniform draws the common case needs (first voter succeeds).  Values go
        through a FIFO, so the draws reach the scores in the same order whether or not they were drawn ahead.
        (Parity with a same-seed run of the reference holds for the FIRST vote only: between rounds the reference
        also draws from the global torch generator whenever a DataLoader iterator is created and when
        ModelVerifier builds a fresh model, which this engine does not replay.  The jitter is a +-1e-4 tie-break,
        so only elections between near-equal scores can differ.)"""
        if self._prefetched is not None:
            return
        n_sel = max(1, int(self.cfg.num_participants * self.N))
        sel = random.sample(range(self.N), n_sel)
        if self.cfg.aggregation != "gossip":          # gossip has no vote: no tie-break draws
            need = (n_sel - 1) + (n_sel if self.update_type == "mse_avg" else 0)
            while len(self._rand_fifo) < need:
                self._rand_fifo.append(torch.rand(1).item())
        self._prefetched = (round_idx, sel)

    def make_control(self, round_idx: int) -> RoundControl:
        cfg = self.cfg
        n_sel = max(1, int(cfg.num_participants * self.N))
        if self._prefetched is not None:
            _, sel = self._prefetched
            self._prefetched = None
        else:
            sel = random.sample(range(self.N), n_sel)
        if cfg.aggregation == "gossip":               # no election: every selected client merges its neighbourhood
            return RoundControl(round_idx, sel, -1, [1.0] * n_sel, sel[0], -1)
        quota = cfg.max_aggregation_threshold
        voter_pos, jitter = -1, [1.0] * n_sel
        for k in range(n_sel):
            draws = [1.0 + (self._next_rand() - 0.5) * 0.0002 for _ in range(n_sel - 1)]
            if any(self.agg_count_host[g] < quota for p, g in enumerate(sel) if p != k):
                voter_pos = k
                it = iter(draws)
                jitter = [1.0 if p == k else next(it) for p in range(n_sel)]
                break
        if voter_pos >= 0 and self.update_type == "mse_avg":
            for _ in range(n_sel):          # quirk Q7: aggregate_models draws once per client
                self._next_rand()
        verify_client = self.N - 1 if (cfg.reference_compat and cfg.verification_method == "val") else -1
        return RoundControl(round=round_idx, sel=sel, voter_pos=voter_pos, jitter=jitter,
                            vote_client=sel[0], verify_client=verify_client)

    def _voter_is_decided(self, next_sel: Sequence[int], cur_sel: Sequence[int]) -> bool:
        """Can the next round's voter be determined before this round's aggregator is known?  The only host
        state the control block depends on is the aggregation count, and this round raises it by one for one
        client of ``cur_sel``.  Scan the candidates like ``make_control``: a candidate is surely valid if some
        other selected client stays under the quota even after such an increment, surely invalid if none is
        under it now; anything in between has to wait for the read-back."""
        quota = self.cfg.max_aggregation_threshold
        cur = set(cur_sel)
        for k in range(len(next_sel)):
            low = [g for p, g in enumerate(next_sel) if p != k and self.agg_count_host[g] < quota]
            if any(self.agg_count_host[g] + (1 if g in cur else 0) < quota for g in low):
                return True
            if low:
                return False
        return True                       # nobody can vote, whatever this round elects

    def _sel_local(self, sel: Sequence[int]) -> List[int]:
        return [g - self.first for g in sel if self.first <= g < self.first + self.n_local]

    # ------------------------------------------------------------------ one round
    def run_round(self, round_idx: int, ctl: Optional[RoundControl] = None,
                  evaluate: bool = True, want_latents: Optional[bool] = None,
                  prefetch_next: bool = False, pipeline_next: bool = False) -> RoundResult:
        """One federated round.  ``prefetch_next`` prepares the next round's selection and random draws on
        the host while the device executes this round (use it when more rounds follow).

        ``pipeline_next`` is a promise that ``run_round(round_idx + 1)`` WILL be called next with default
        control: this call then also uploads the next round's control block and enqueues its local training
        before it waits for its own results, so the device never idles on the host's bookkeeping (staged
        data, if any, is committed first).  The next call picks up from there.  Results are identical to
        the unpipelined sequence; if the next voter cannot be determined yet (quota nearly exhausted) the
        call silently falls back to the plain order."""
        self._prefetch_next = (prefetch_next or pipeline_next) and ctl is None
        self._pipeline_next = pipeline_next and self._prefetch_next
        be, st, sh, cm, tm = self.backend, self.state, self.sh, self.comm, self.timer
        pre, self._pre = self._pre, None
        if pre is not None:
            if pre[0] != round_idx or ctl is not None:
                raise RuntimeError(f"round {pre[0]} was already started by run_round(..., pipeline_next=True); "
                                   "it has to be the next round, with the engine's own control")
            ctl = pre[1]                  # control uploaded, training enqueued
        else:
            ctl = ctl or self.make_control(round_idx)
            if hasattr(be, "begin_round"):
                be.begin_round(ctl)
            with tm.phase("train"):
                be.train(st, sh, self._sel_local(ctl.sel), self.hp)
        agg_pos = None
        if self.cfg.aggregation == "gossip":
            self._gossip_phase(ctl)
        elif ctl.voter_pos >= 0:
            with tm.phase("vote"):
                scores = be.vote_scores(st, sh, ctl, self.spec)
                if not getattr(be, "exchanges_scores", False):     # CUDA multi-GPU exchanges in-kernel (P2P boards)
                    scores = cm.sum_(scores)
            # The election waits for every rank's vote scores, i.e. for the slowest rank's local training.  The
            # merge weights do not depend on who is elected, so their scoring (the heavy part of the merge) is
            # enqueued first and the cross-rank wait hides under it.
            early_scores = getattr(be, "exchanges_scores", False) and self.update_type == "mse_avg"
            mse = None
            if early_scores:
                with tm.phase("score"):
                    mse = be.merge_scores(st, sh, ctl)
            with tm.phase("elect"):
                agg_pos = be.elect(st, scores, ctl, self.spec)
            with tm.phase("merge"):
                if early_scores:
                    pass
                elif self.update_type == "mse_avg":
                    mse = be.merge_scores(st, sh, ctl)
                    if not getattr(be, "exchanges_scores", False):
                        mse = cm.sum_(mse)
                elif self.update_type == "fusion_avg":
                    mse = be.fusion_scores(st, sh, ctl, self.cfg.fusion_rows)
                    if not getattr(be, "exchanges_scores", False):
                        mse = cm.sum_(mse)
                rows = None
                if cm.is_dist and not getattr(be, "reads_peers", False):
                    rows = cm.gather_rows(st.params, ctl.sel, self.first)
                theta = be.merge(st, sh, ctl, self.spec, mse, agg_pos, rows)
            with tm.phase("verify"):
                be.verify_update(st, sh, ctl, self.spec, theta, agg_pos)
        return self._finish_round(ctl, agg_pos, evaluate, want_latents)

    def _finish_round(self, ctl: RoundControl, agg_pos, evaluate: bool,
                      want_latents: Optional[bool]) -> RoundResult:
        be, st, sh, cm, tm, cfg = self.backend, self.state, self.sh, self.comm, self.timer, self.cfg
        if want_latents is None:
            want_latents = cfg.write_latents and self.model_type == "hybrid"
        metrics: List[float] = []
        lat, auc64, auc_dev = None, None, None
        if getattr(self, "_prefetch_next", False):
            self.prefetch_control(ctl.round + 1)       # host work, overlaps with the kernels enqueued so far
        pipelined = (getattr(self, "_pipeline_next", False) and hasattr(be, "readback_async")
                     and (cfg.aggregation == "gossip" or self._voter_is_decided(self._prefetched[1], ctl.sel)))
        # pipelined rounds evaluate on a side stream from a parameter snapshot, concurrently with the next
        # round's local training (which leaves more than half of the SMs idle)
        side_eval = (pipelined and evaluate and not want_latents and cfg.metric == "AUC"
                     and hasattr(be, "evaluate_side"))
        if evaluate:
            with tm.phase("evaluate"):
                if side_eval:
                    be.evaluate_side(st, sh, self.model_type)
                else:
                    auc_dev, auc64, lat = be.evaluate(st, sh, self.model_type, want_latents, cfg.metric)
        # ---- single readback of the round's results (global lists) ------------------------
        if hasattr(be, "readback") and (not evaluate or auc64 is None):
            handle = be.readback_async(st, agg_pos, evaluate, side_eval) if side_eval else \
                be.readback_async(st, agg_pos, evaluate)
            if pipelined:
                # the next round's upload and training go in behind this round's work, before the host blocks
                self.commit_staged_data()
                nxt = self.make_control(ctl.round + 1)
                be.begin_round(nxt)
                with tm.phase("train"):
                    be.train(st, sh, self._sel_local(nxt.sel), self.hp)
                self._pre = (nxt.round, nxt)
            a, epochs_all, rejected, auc_all = be.readback_wait(handle)
            if evaluate:
                metrics = auc_all
        else:
            a = int(agg_pos[0]) if agg_pos is not None else -1
            if evaluate:
                if auc64 is None:
                    auc64 = [float(x) for x in auc_dev.tolist()]
                metrics = self._gather_list(auc64)
            rejected = self._gather_list(st.rejected.tolist())
            epochs_all = self._gather_list(st.epochs_run.tolist())
        aggregator = ctl.sel[a] if a >= 0 else -1
        if aggregator >= 0:
            self.agg_count_host[aggregator] += 1
        verification = []
        if aggregator >= 0:
            for g in range(self.N):
                if g != aggregator:
                    verification.append({"client_id": g, "rejected_updates": int(rejected[g]),
                                         "is_verified": int(rejected[g]) == 0})
        ep = {g: int(epochs_all[g]) for g in ctl.sel}
        samples = 0
        for g in self._sel_global_local(ctl.sel):
            c = g - self.first
            samples += ep[g] * (sh.train.count(c) + sh.valid.count(g))
        if want_latents and lat is not None:
            self._stash_latents(ctl.round, lat)
        res = RoundResult(ctl.round, list(ctl.sel), aggregator, metrics, verification, ep, samples)
        if evaluate:
            self.last_metrics = metrics
        self.history.append(res)
        return res

    def _sel_global_local(self, sel):
        return [g for g in sel if self.first <= g < self.first + self.n_local]

    def _gather_list(self, local: Sequence) -> List:
        if not self.comm.is_dist:
            return list(local)
        parts = self.comm.all_gather_object(list(local))
        out: List = []
        for p in parts:
            out.extend(p)
        return out

    def _stash_latents(self, round_idx: int, lat: List[np.ndarray]) -> None:
        labels = [self.sh.test.labels(c).detach().cpu().numpy().astype("float32")
                  for c in range(self.n_local)]
        local = {self.sh.names[c]: (lat[c], labels[c]) for c in range(self.n_local)}
        merged: Dict[str, Tuple[np.ndarray, np.ndarray]] = {}
        for part in self.comm.all_gather_object(local):
            merged.update(part)
        self.latents[round_idx] = merged

    # ------------------------------------------------------------------ gossip round
    def neighbours(self, round_idx: int) -> List[List[int]]:
        """Host copy of the round's gossip graph (the CUDA backend generates the same graph in its kernel)."""
        cfg = self.cfg
        if cfg.topology.startswith("random") or self._neigh_cache is None:
            self._neigh_cache = topology.build(cfg.topology, self.N, cfg.topology_k,
                                               cfg.topology_seed, round_idx)
        return self._neigh_cache

    def _gossip_phase(self, ctl: RoundControl) -> None:
        be, st, sh, cm, tm, cfg = self.backend, self.state, self.sh, self.comm, self.timer, self.cfg
        with tm.phase("gossip"):
            be_gossip = getattr(be, "gossip", None)
            if be_gossip is None:
                raise RuntimeError(f"backend {be.name} has no gossip op")
            if getattr(be, "device_topology", False):
                be_gossip(st, sh, ctl, self.spec, None, None,
                          topo=(cfg.topology, cfg.topology_k, cfg.topology_seed))
            else:
                rows = None
                if cm.is_dist and not getattr(be, "reads_peers", False):
                    rows = cm.gather_rows(st.params, list(range(self.N)), self.first)
                be_gossip(st, sh, ctl, self.spec, self.neighbours(ctl.round), rows)

    # ------------------------------------------------------------------ full experiment for one combination
    def fit(self, stop: Optional[GlobalStop] = None, write: bool = True) -> List[RoundResult]:
        cfg = self.cfg
        stop = stop or GlobalStop(cfg.global_patience)
        root = cfg.output_root
        res_file = artefacts.results_file(root, cfg.network_size, cfg.no_Exp, self.run, cfg.metric,
                                          cfg.scen_name, cfg.num_participants, self.model_type,
                                          self.update_type)
        ver_file = f"{root}/{cfg.checkpoint_dir}/Run_{self.run}/verification_results.json"
        out = []
        ckpt = f"{root}/{cfg.checkpoint_dir}/Run_{self.run}/resume_{self.model_type}_{self.update_type}.pt"
        r0 = 0
        if cfg.resume and os.path.exists(ckpt if self.comm.world == 1 else f"{ckpt}.rank{self.comm.rank}"):
            r0 = self.load_checkpoint(ckpt)
            stop.min_val_loss, stop.worse = self.checkpoint_extra.get("stop", (stop.min_val_loss, stop.worse))
            log.info("Resumed %s + %s run %d at round %d from %s", self.model_type, self.update_type, self.run, r0 + 1, ckpt)
            if self.checkpoint_extra.get("stopped"):
                r0 = cfg.num_rounds                      # this combination had already stopped early
        for r in range(r0, cfg.num_rounds):
            log.info("Starting round %d/%d", r + 1, cfg.num_rounds)
            res = self.run_round(r, prefetch_next=r + 1 < cfg.num_rounds)
            out.append(res)
            if res.aggregator >= 0:
                log.info("Client %d selected as aggregator", res.aggregator + 1)
            else:
                log.warning("No aggregator selected for this round")
            for i, m in enumerate(res.client_metrics):
                log.info("Client %d %s score: %s", i + 1, cfg.metric, m)
            if write:
                if cfg.save_client_models:
                    self.save_client_models(res)
                if self.comm.rank == 0:
                    if res.aggregator >= 0:
                        artefacts.append_verification(ver_file, r + 1, res.verification)
                    artefacts.append_round_metrics(res_file, r + 1, res.client_metrics,
                                                   self.update_type, self.model_type)
            stopped = stop.update(res.client_metrics)
            if cfg.checkpoint_every > 0 and ((r + 1) % cfg.checkpoint_every == 0 or stopped or r + 1 == cfg.num_rounds):
                self.save_checkpoint(ckpt, extra={"stop": (stop.min_val_loss, stop.worse), "stopped": bool(stopped)})
            if stopped:
                log.info("Early stopping in global round!")
                break
        if write and self.comm.rank == 0 and self.latents:
            artefacts.write_latents(
                artefacts.latent_file(root, cfg.network_size, cfg.no_Exp, self.run, self.model_type,
                                      self.update_type), self.latents)
        return out

    def final_best(self) -> float:
        """Max client metric after the last round (``src/main.py:368-374``, quirk Q18)."""
        if self.last_metrics is None:
            _, auc64, _ = self.backend.evaluate(self.state, self.sh, self.model_type, False, self.cfg.metric)
            self.last_metrics = self._gather_list(auc64)
        return max(self.last_metrics)

    def save_client_models(self, res: RoundResult) -> None:
        """Write ``model.cpt`` (best-validation weights of this round's run) and
        ``training_tracking.pkl`` for every locally-owned client that trained."""
        cfg, st = self.cfg, self.state
        sel_local = self._sel_local(res.selected)
        if not sel_local:
            return
        idx = torch.tensor(sel_local, dtype=torch.long, device=st.device)
        best = st.best[idx].cpu()
        logs = st.loss_log[idx].cpu()
        for k, c in enumerate(sel_local):
            g = self.first + c
            d = artefacts.client_model_dir(cfg.output_root, cfg.network_size, cfg.no_Exp, self.run,
                                           cfg.scen_name, self.model_type, self.update_type,
                                           self.names_global[g])
            artefacts.save_client_model(d, self.layout.unpack(best[k]))
            n = res.epochs_run.get(g, 0)
            artefacts.save_tracking(d, [(float(logs[k, e, 0]), float(logs[k, e, 1])) for e in range(n)])

    # ------------------------------------------------------------------ streaming data
    def update_local_data(self, train_x: Optional[torch.Tensor] = None, valid_x: Optional[torch.Tensor] = None
                          ) -> None:
        """Replace the rank's local training / validation rows with fresh traffic (same shapes), e.g. from
        pinned host memory each round; copies are asynchronous on the current stream."""
        sh = self.sh
        touched = []
        if train_x is not None:
            sh.train.x.copy_(train_x, non_blocking=True)
            touched.append(sh.train.x)
        if valid_x is not None:
            v0, v1 = sh.valid.off_host[self.first], sh.valid.off_host[self.first + self.n_local]
            sh.valid.x[v0:v1].copy_(valid_x, non_blocking=True)
            touched.append(sh.valid.x)
        if touched and hasattr(self.backend, "data_updated"):
            self.backend.data_updated(*touched)

    def stage_next_data(self, train_x: Optional[torch.Tensor] = None, valid_x: Optional[torch.Tensor] = None) -> None:
        """Start uploading the NEXT round's local rows (e.g. from pinned host memory) into shadow shards on
        a side stream, so the copy overlaps with the round that is about to run on the current shards.
        Call ``commit_staged_data()`` before the round that should see them."""
        sh, dev = self.sh, self.state.device
        if dev.type != "cuda":
            self._staged = (train_x, valid_x, None)
            return
        if getattr(self, "_copy_stream", None) is None:
            self._copy_stream = torch.cuda.Stream(device=dev)
            self._shadow_train = sh.train.x.clone()
            self._shadow_valid = sh.valid.x.clone()
            torch.cuda.current_stream(dev).synchronize()
        v0, v1 = sh.valid.off_host[self.first], sh.valid.off_host[self.first + self.n_local]
        with torch.cuda.stream(self._copy_stream):
            if train_x is not None:
                self._shadow_train.copy_(train_x, non_blocking=True)
            if valid_x is not None:
                self._shadow_valid[v0:v1].copy_(valid_x, non_blocking=True)
            ev = torch.cuda.Event()
            ev.record(self._copy_stream)
        self._staged = (train_x is not None, valid_x is not None, ev)

    def commit_staged_data(self) -> None:
        """Make the staged rows current (pointer swap after the side-stream copy has finished)."""
        staged = getattr(self, "_staged", None)
        if staged is None:
            return
        self._staged = None
        sh, dev = self.sh, self.state.device
        if dev.type != "cuda":
            self.update_local_data(staged[0], staged[1])
            return
        has_t, has_v, ev = staged
        torch.cuda.current_stream(dev).wait_event(ev)
        touched = []
        if has_t:
            sh.train.x, self._shadow_train = self._shadow_train, sh.train.x
            touched.append(sh.train.x)
        if has_v:
            sh.valid.x, self._shadow_valid = self._shadow_valid, sh.valid.x
            touched.append(sh.valid.x)
        if touched and hasattr(self.backend, "data_updated"):
            self.backend.data_updated(*touched)

    def close(self) -> None:
        """Free backend resources that outlive python garbage collection (peer-mapped slabs).  Collective when
        distributed.  The engine is unusable afterwards."""
        if self._pre is not None:
            self.backend.synchronize()
            self._pre = None
        be_close = getattr(self.backend, "close", None)
        if be_close is not None:
            self.state.params = None          # aliases the parameter slab
            be_close()

    # ------------------------------------------------------------------ resume (absent in the reference)
    def snapshot(self) -> Dict:
        if self._pre is not None:
            raise RuntimeError("a pipelined round is in flight (run_round(..., pipeline_next=True)); finish it first")
        return {"state": self.state.state_dict(), "agg_count_host": list(self.agg_count_host),
                "py_random": random.getstate(), "np_random": np.random.get_state(),
                "torch_random": torch.get_rng_state(), "rounds_done": getattr(self, "_rounds_offset", 0) + len(self.history),
                "model_type": self.model_type, "update_type": self.update_type, "run": self.run,
                "rand_fifo": list(self._rand_fifo), "prefetched": self._prefetched}

    def restore(self, snap: Dict) -> None:
        self.state.load_state_dict(snap["state"])
        self.agg_count_host = list(snap["agg_count_host"])
        random.setstate(snap["py_random"])
        np.random.set_state(snap["np_random"])
        torch.set_rng_state(snap["torch_random"])
        self._rand_fifo = list(snap.get("rand_fifo", []))
        self._prefetched = snap.get("prefetched")
        self._rounds_offset = int(snap.get("rounds_done", 0)) - len(self.history)

    def save_checkpoint(self, path: str, extra: Optional[Dict] = None) -> str:
        """``snapshot()`` on disk: parameters, Adam state, verifier history, quotas, RNG streams, rounds done.  The
        reference never reloads anything; ``load_checkpoint`` on an engine built from the same config
        and data continues the experiment with the rounds it would have run.  On several ranks every rank writes its own
        file (``path`` + ``.rank<r>``): a rank's state holds only the clients it owns."""
        snap = self.snapshot()
        snap["format"] = 1
        snap["n_clients"] = int(self.state.n_clients)
        snap["extra"] = dict(extra or {})
        if self.comm is not None and getattr(self.comm, "world", 1) > 1:
            path = f"{path}.rank{self.comm.rank}"
        os.makedirs(os.path.dirname(os.path.abspath(path)), exist_ok=True)
        tmp = path + ".tmp"
        torch.save(snap, tmp)
        os.replace(tmp, path)                      # never leave a half-written checkpoint behind
        return path

    def load_checkpoint(self, path: str) -> int:
        """Restore a ``save_checkpoint`` file; returns the index of the next round to run."""
        if self.comm is not None and getattr(self.comm, "world", 1) > 1:
            path = f"{path}.rank{self.comm.rank}"
        snap = torch.load(path, map_location="cpu", weights_only=False)
        if snap.get("format") != 1:
            raise ValueError(f"{path}: not a checkpoint of this engine")
        if (snap["model_type"], snap["update_type"]) != (self.model_type, self.update_type) or \
                snap["n_clients"] != int(self.state.n_clients):
            raise ValueError(f"{path}: written for {snap['model_type']} + {snap['update_type']} with {snap['n_clients']} "
                             f"local clients, this engine is {self.model_type} + {self.update_type} with "
                             f"{int(self.state.n_clients)}")
        self.restore(snap)
        self.checkpoint_extra = dict(snap.get("extra", {}))
        return int(snap["rounds_done"])
