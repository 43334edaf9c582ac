"""CUDA backend: maps each round-engine operation to the sm_100a kernels of ``csrc/``.

Kernel launches per round (elected mode): 1 train + (1 normalise + 1 score + 1
reduce + 1 elect) for the vote + (1 score + 1 reduce) for MSE weights + 1 merge +
(1 score + 1 verify) + (1-2 score + 1 AUC) for evaluation; the host uploads one
small control block and reads one small result block.  The reference issues
O(clients x steps x 40) launches and a host sync per step for the same round
(SURVEY.md §3.2, §7.1).
"""
from __future__ import annotations

import types
import ctypes as C
from typing import TYPE_CHECKING, Dict, List, Optional, Sequence

import numpy as np
import torch

if TYPE_CHECKING:  # avoid the ops <-> trainer import cycle
    from ..data.shards import ShardStore
    from ..trainer.state import FederationState

from ..layout import D_PAD, ParamLayout
from . import _lib
from .backend import MergeSpec, RoundControl
from .oracle import TrainHP

TILE = 64


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def make_tiles(segments: Sequence[Sequence[int]], tile: int = TILE) -> torch.Tensor:
    """Tile table ``[n, 4] = {slot, row0, nrows, out0}`` covering ``(slot, row0, nrows, out0)`` segments."""
    out = []
    for slot, row0, nrows, out0 in segments:
        for s in range(0, nrows, tile):
            out.append((slot, row0 + s, min(tile, nrows - s), out0 + s))
    return torch.tensor(out, dtype=torch.int32).reshape(-1, 4)


class CudaBackend:
    name = "cuda"
    reads_peers = True   # merge / gossip kernels read other ranks' slabs themselves

    def __init__(self, layout: ParamLayout, device, score_precision: str = "fp32"):
        if score_precision not in ("fp32", "tf32", "tf32x3", "fp16x3"):
            raise ValueError("score_precision must be 'fp32', 'tf32', 'tf32x3' or 'fp16x3'")
        # scoring forwards of merge / verify / evaluate: exact FFMA kernel, or the tcgen05 kernel (single
        # TF32 pass or error-compensated 3xTF32).  Vote scoring always uses the exact kernel: the election
        # is an argmin over nearly equal scores.
        self.score_precision = score_precision
        # "fp16x3": the MSE-weight scoring of the merge runs on the fused fp16 hi/lo kernel (csrc/score_fused_f16.cu,
        # fp32-grade like 3xTF32 at twice the tensor rate); verification / evaluation forwards stay on 3xTF32.
        self.tc = {"fp32": 0, "tf32": 1, "tf32x3": 3, "fp16x3": 3}[score_precision]
        self.fused_f16 = score_precision == "fp16x3"
        self.tile = 128 if self.tc else TILE
        if (layout.dim, layout.hidden, layout.latent) != (115, 27, 7):
            raise ValueError("the sm_100a kernels are specialised for the 115-27-7 autoencoder; "
                             "use backend='oracle' for other shapes")
        self.layout = layout
        self.device = torch.device(device)
        if self.device.type != "cuda":
            raise RuntimeError("CudaBackend needs a CUDA device")
        self.lib = _lib.load()
        self.launches = 0
        self._slabs = []
        self._tile_cache: Dict = {}
        self._resid: Dict = {}

    # ------------------------------------------------------------------ setup
    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    def alloc_param_slab(self, n_local: int, padded: int, comm) -> torch.Tensor:
        from ..parallel.symmetric import SymmetricSlab
        with torch.cuda.device(self.device):
            n_max = n_local
            if comm.is_dist:
                n_max = int(max(comm.all_gather_object(n_local)))
            self.param_slab = SymmetricSlab(n_local, padded, comm, rows_max=n_max)
            self.snap_slab = None
        self._slabs.append(self.param_slab)
        return self.param_slab.tensor

    def bind(self, st: FederationState, sh: ShardStore, comm, hp: TrainHP, spec: MergeSpec, model_type: str):
        dev = self.device
        self.st, self.sh, self.comm = st, sh, comm
        f32 = dict(dtype=torch.float32, device=dev)
        i32 = dict(dtype=torch.int32, device=dev)
        n_sel_max = max(1, sh.n_global)
        M = self._M = n_sel_max
        # one control block per round: [sel | jitter (fp32 bits) | sel_local], uploaded in a single copy
        # from a small ring of pinned staging buffers (a slot is reused only after its copy completed)
        self.ctl_dev = torch.zeros(3 * M, **i32)
        self.ctl_sel = self.ctl_dev[0:M]
        self.ctl_jit = self.ctl_dev[M:2 * M].view(torch.float32)
        self.ctl_sel_local = self.ctl_dev[2 * M:3 * M]
        self._pin = [torch.zeros(3 * M, dtype=torch.int32).pin_memory() for _ in range(4)]
        self._pin_ev = [None] * 4
        self._pin_i = 0
        self.n_sel_local = 0
        self.agg_pos = torch.full((2,), -1, **i32)
        self.scores = torch.zeros(n_sel_max, **f32)
        self.mse = torch.zeros(n_sel_max, **f32)
        self.weights = torch.zeros(n_sel_max, **f32)
        self.theta = torch.zeros(self.layout.padded, **f32)
        max_valid = sh.valid.max_count()
        self.vote_norm = torch.zeros(max(1, max_valid), D_PAD, **f32)        # scratch: normalised vote matrix
        self.vote_sse = torch.zeros(n_sel_max * max(1, max_valid), **f32)    # scratch
        self.dev_rows = sh.dev.shape[0]
        self.merge_sse = None                                                 # lazily [n_sel * dev_rows]
        self.verify_sse = torch.zeros(max(1, sh.valid.x.shape[0]), **f32)
        self.test_sse = torch.zeros(max(1, sh.test.x.shape[0]), **f32)
        self.test_lat = torch.zeros(max(1, sh.test.x.shape[0]), 8, **f32)
        self.train_lat = torch.zeros(max(1, sh.train.x.shape[0]), 8, **f32)
        self.auc = torch.zeros(max(1, sh.n_clients), dtype=torch.float64, device=dev)
        self.scores_out = torch.zeros(max(1, sh.test.x.shape[0]), **f32)
        # static tile tables
        C_ = sh.n_clients
        self.tiles_test = make_tiles([(c, sh.test.off_host[c], sh.test.count(c), sh.test.off_host[c])
                                      for c in range(C_)], self.tile).to(dev)
        self.tiles_train = make_tiles([(c, sh.train.off_host[c], sh.train.count(c), sh.train.off_host[c])
                                       for c in range(C_)], self.tile).to(dev)
        f = sh.first_client
        self.tiles_verify_own = make_tiles([(0, sh.valid.off_host[f + c], sh.valid.count(f + c),
                                             sh.valid.off_host[f + c]) for c in range(C_)], self.tile).to(dev)
        self.verify_off_own = torch.tensor([sh.valid.off_host[f + c] for c in range(C_ + 1)], **i32)
        self.peer_table = self.param_slab.table
        # multi-GPU: peer-mapped control block (score boards + arrival counters) -> no NCCL on the merge path
        self.exchanges_scores = comm.is_dist
        self.seq = [0, 0, 0]
        self.nvls = None
        if comm.is_dist:
            from ..parallel.symmetric import SymmetricSlab
            if n_sel_max > _lib.MAX_SEL:
                raise ValueError(f"{n_sel_max} selected clients exceed FM_MAX_SEL={_lib.MAX_SEL}")
            with torch.cuda.device(dev):
                self.ctl_slab = SymmetricSlab(1, _lib.CTL_FLOATS, comm)
            self._slabs.append(self.ctl_slab)
            self.ctl = self.ctl_slab.tensor.view(-1)
            self.ctl_arrive_ptr = self.ctl_slab.ptr + 4 * 2 * _lib.MAX_SEL
            self.done_counter = torch.zeros(1, dtype=torch.int32, device=dev)
            # NVLS: [partial | theta] in one symmetric buffer with a multicast view; theta then IS that buffer's
            # second half, so verification / scoring read the switch-broadcast result in place
            from ..parallel import nvls
            self.nvls = nvls.allocate(comm, 2 * self.layout.padded)
            if self.nvls is not None:
                self.theta = self.nvls.tensor[self.layout.padded:]
                self._nvls_counter = torch.zeros(2, dtype=torch.int32, device=dev)
                self._nvls_seq = 0

    def begin_round(self, ctl: RoundControl) -> None:
        """Upload the round's control block: selection (global + local ids) and jitter."""
        n, M, sh = len(ctl.sel), self._M, self.sh
        i = self._pin_i
        self._pin_i = (i + 1) % len(self._pin)
        if self._pin_ev[i] is not None:
            self._pin_ev[i].synchronize()
        pin = self._pin[i]
        pin[0:n] = torch.tensor(ctl.sel, dtype=torch.int32)
        pin[M:M + n].view(torch.float32)[:] = torch.tensor(ctl.jitter, dtype=torch.float32)
        loc = [g - sh.first_client for g in ctl.sel if sh.first_client <= g < sh.first_client + sh.n_clients]
        self.n_sel_local = len(loc)
        if loc:
            pin[2 * M:2 * M + len(loc)] = torch.tensor(loc, dtype=torch.int32)
        self.ctl_dev.copy_(pin, non_blocking=True)
        ev = torch.cuda.Event()
        ev.record()
        self._pin_ev[i] = ev
        self.agg_pos.fill_(-1)

    def _residual(self, X: torch.Tensor, refresh: bool = False) -> torch.Tensor:
        """Residual shard ``X - trunc_tf32(X)`` of a data matrix for the 3xTF32 scoring kernel (made once
        per matrix; call with ``refresh=True`` after the matrix was overwritten with new rows)."""
        key = X.data_ptr()
        ent = self._resid.get(key)
        if ent is None or ent.shape != X.shape:
            ent = torch.empty_like(X)
            self._resid[key] = ent
            refresh = True
        if refresh:
            _lib.check(self.lib.fm_split_residual(_ptr(X), _ptr(ent), X.shape[0], self._stream()), "split_residual")
            self.launches += 1
        return ent

    def data_updated(self, *mats: torch.Tensor) -> None:
        """Tell the backend that device-resident data matrices were overwritten (fresh traffic)."""
        if self.tc == 3:
            for X in mats:
                if X.data_ptr() in self._resid:
                    self._residual(X, refresh=True)

    def _tiles(self, key, builder) -> torch.Tensor:
        t = self._tile_cache.get(key)
        if t is None:
            t = builder().to(self.device)
            self._tile_cache[key] = t
        return t

    def _score(self, slab: torch.Tensor, n_local: int, first: int, slot_client: Optional[torch.Tensor],
               X: torch.Tensor, tiles: torch.Tensor, out_sse: Optional[torch.Tensor],
               out_lat: Optional[torch.Tensor]) -> None:
        a = _lib.FmScoreArgs(_ptr(slab), _ptr(slot_client), first, n_local, _ptr(X), _ptr(tiles),
                             tiles.shape[0], _ptr(out_sse), _ptr(out_lat))
        if self.tc:
            lo = self._residual(X) if self.tc == 3 else None
            _lib.check(self.lib.fm_score_rows_tc(C.byref(a), X.shape[0], self.tc, _ptr(lo), self._stream()),
                       "score_rows_tc")
        else:
            _lib.check(self.lib.fm_score_rows(C.byref(a), self._stream()), "score_rows")
        self.launches += 1

    # ------------------------------------------------------------------ T1
    def train(self, st: FederationState, sh: ShardStore, sel_local: Sequence[int], hp: TrainHP) -> None:
        n = len(sel_local)
        assert n == self.n_sel_local, "begin_round() must upload the control block before train()"
        st.epochs_run.zero_()
        if n == 0:
            return
        if hp.batch_size > 32:      # thesis setting (batch 64): beyond the shared-memory budget of the persistent kernel
            from .generic_backend import generic_train
            generic_train(self.lib, self.layout, st, sh, self.ctl_sel_local, n, hp, self._stream(), self)
            self.launches += 1
            return
        a = _lib.FmTrainArgs(
            _ptr(st.params), _ptr(st.adam_m), _ptr(st.adam_v), _ptr(st.adam_t), _ptr(st.anchor), _ptr(st.best),
            _ptr(st.loss_log), _ptr(st.epochs_run), _ptr(st.trained_rounds),
            _ptr(sh.train.x), _ptr(sh.train.off), _ptr(sh.valid.x), _ptr(sh.valid.off),
            _ptr(self.ctl_sel_local), n, sh.first_client, hp.epochs, hp.batch_size, hp.patience,
            st.loss_log.shape[1], hp.lr, hp.shrink_lambda, hp.fedprox_mu,
            _ptr(getattr(self, "phase_cycles", None)))
        _lib.check(self.lib.fm_train_clients(C.byref(a), self._stream()), "train_clients")
        self.launches += 1

    # ------------------------------------------------------------------ T2 vote
    def vote_scores(self, st, sh, ctl: RoundControl, spec: MergeSpec) -> torch.Tensor:
        n = len(ctl.sel)
        v0, rows = sh.valid.off_host[ctl.vote_client], sh.valid.count(ctl.vote_client)
        V = sh.valid.x[v0:v0 + rows]
        tiles = self._tiles(("vote", n, rows), lambda: make_tiles([(p, 0, rows, p * rows) for p in range(n)]))
        a = _lib.FmVoteArgs(_ptr(st.params), _ptr(self.ctl_sel), n, sh.first_client, st.n_clients, _ptr(V), rows,
                            spec.vote_chunk, _ptr(self.ctl_jit), ctl.voter_pos, _ptr(tiles), tiles.shape[0],
                            _ptr(self.vote_norm), _ptr(self.vote_sse), _ptr(self.scores))
        _lib.check(self.lib.fm_vote_scores(C.byref(a), self._stream()), "vote_scores")
        self.launches += 3
        if self.exchanges_scores:
            return self._publish(self.scores, n, 0)
        return self.scores[:n]

    def _publish(self, local: torch.Tensor, n: int, which: int) -> torch.Tensor:
        """Write this rank's entries into every rank's board `which` and signal arrival; returns the view of
        the local board that is complete once `arrive[which] >= world * seq[which]` (consumers wait in-kernel)."""
        sh, st = self.sh, self.st
        _lib.check(self.lib.fm_publish(_ptr(local), _ptr(self.ctl_sel), sh.first_client, st.n_clients, n,
                                       _ptr(self.ctl_slab.table), self.comm.world, which, self._stream()), "publish")
        self.launches += 1
        self.seq[which] += 1
        return self.ctl[which * _lib.MAX_SEL: which * _lib.MAX_SEL + n]

    def _arrive(self, which: int):
        if not self.exchanges_scores:
            return 0, 0
        return self.ctl_arrive_ptr + 4 * which, self.comm.world * self.seq[which]

    def elect(self, st, scores: torch.Tensor, ctl: RoundControl, spec: MergeSpec) -> torch.Tensor:
        arrive, expect = self._arrive(0)
        _lib.check(self.lib.fm_elect(_ptr(scores), _ptr(self.ctl_sel), len(ctl.sel), ctl.voter_pos,
                                     _ptr(st.agg_count), spec.max_aggregations, _ptr(self.agg_pos),
                                     arrive, expect, self._stream()), "elect")
        self.launches += 1
        return self.agg_pos

    # ------------------------------------------------------------------ T2 merge
    def _merge_scores_fused(self, st, sh, ctl: RoundControl) -> torch.Tensor:
        """prep + ONE kernel: fp16 hi/lo tcgen05 forward of every locally owned selected model over the dev matrix,
        per-model MSE finished in the kernel (no SSE matrix, no chunk_mean)."""
        n, R = len(ctl.sel), self.dev_rows
        if getattr(self, "_images", None) is None:
            self._images = torch.empty(max(1, st.n_clients) * _lib.TC_IMAGE_BYTES, dtype=torch.uint8,
                                       device=self.device)
            self._owned = torch.zeros(max(1, sh.n_global) + 1, dtype=torch.int32, device=self.device)
        if getattr(self, "_fused_partial", None) is None:
            need = int(self.lib.fm_score_fused_scratch_floats(R, max(1, st.n_clients)))
            self._fused_partial = torch.empty(need, dtype=torch.float32, device=self.device)
            self._fused_gbar = torch.zeros(2, dtype=torch.int32, device=self.device)
        self.mse.zero_()
        a = _lib.FmFusedScoreArgs(_ptr(st.params), _ptr(self.ctl_sel), n, sh.first_client, st.n_clients,
                                  _ptr(sh.dev), R, _ptr(self._images), _ptr(self._owned[1:]), _ptr(self._owned[:1]),
                                  _ptr(self._fused_partial), _ptr(self.mse), _ptr(self._fused_gbar),
                                  _ptr(getattr(self, "mm_cycles", None)))
        _lib.check(self.lib.fm_score_fused_f16(C.byref(a), self._stream()), "score_fused_f16")
        self.launches += 2
        if self.exchanges_scores:
            return self._publish(self.mse, n, 1)
        return self.mse[:n]

    def merge_scores(self, st, sh, ctl: RoundControl) -> torch.Tensor:
        n, R = len(ctl.sel), self.dev_rows
        if self.fused_f16:
            if getattr(self, "_dev_absmax", None) is None:       # static matrix: checked once
                self._dev_absmax = float(sh.dev.abs().max()) if sh.dev.numel() else 0.0
            if self._dev_absmax <= 1.0e4:                        # fp16 head + remainder cover |x| < 1.3e5 exactly enough
                return self._merge_scores_fused(st, sh, ctl)
        if self.merge_sse is None or self.merge_sse.numel() < n * R:
            self.merge_sse = torch.zeros(n * R, dtype=torch.float32, device=self.device)
        if self.tc:
            # tensor-core path: dev tile resident in smem, all locally owned models looped over it
            if getattr(self, "_images", None) is None:
                self._images = torch.empty(max(1, st.n_clients) * _lib.TC_IMAGE_BYTES, dtype=torch.uint8,
                                           device=self.device)
                self._owned = torch.zeros(max(1, sh.n_global) + 1, dtype=torch.int32, device=self.device)
            # (the kernel splits X into TF32 head + remainder itself while parking it in tensor memory: no residual shard)
            a = _lib.FmMultiScoreArgs(_ptr(st.params), _ptr(self.ctl_sel), n, sh.first_client, st.n_clients,
                                      _ptr(sh.dev), None, R, self.tc, _ptr(self._images),
                                      _ptr(self._owned[1:]), _ptr(self._owned[:1]), _ptr(self.merge_sse),
                                      _ptr(getattr(self, "mm_cycles", None)))
            _lib.check(self.lib.fm_score_multi_tc(C.byref(a), self._stream()), "score_multi_tc")
            self.launches += 2
        else:
            tiles = self._tiles(("dev", n, R), lambda: make_tiles([(p, 0, R, p * R) for p in range(n)], self.tile))
            self._score(st.params, st.n_clients, sh.first_client, self.ctl_sel, sh.dev, tiles, self.merge_sse, None)
        _lib.check(self.lib.fm_chunk_mean(_ptr(self.merge_sse), n, R, 0, 0, _ptr(self.ctl_sel), sh.first_client,
                                          st.n_clients, -1, _ptr(self.mse), self._stream()), "chunk_mean")
        self.launches += 1
        if self.exchanges_scores:
            return self._publish(self.mse, n, 1)
        return self.mse[:n]

    def fusion_scores(self, st, sh, ctl: RoundControl, max_rows: int) -> torch.Tensor:
        """``fusion_avg`` weights (the reference's orphaned KDE / Jensen-Shannon similarity) on the device:
        ``csrc/similarity_kernel.cu`` reconstructs the first ``max_rows`` dev rows with every locally owned selected
        model, takes the Gaussian-KDE self log-density of each reconstruction and its JS distance to that of the rows
        themselves; the distances then ride the same board exchange and merge kernel as the MSE scores.
        (fp64 torch oracle of the same quantity: ``utils.similarity.fusion_distances``.)"""
        n = len(ctl.sel)
        rows = int(min(max_rows, self.dev_rows))
        if rows < 2:
            raise ValueError("fusion_avg needs at least two dev rows")
        key = (n, rows)
        if getattr(self, "_kde_key", None) != key:
            need = int(self.lib.fm_kde_js_scratch_floats(n, rows))
            self._kde_rec = torch.empty(need, dtype=torch.float32, device=self.device)
            self._kde_dens = torch.empty((n + 1) * rows, dtype=torch.float32, device=self.device)
            self._kde_key = key
        self.mse.zero_()
        a = _lib.FmKdeJsArgs(_ptr(st.params), _ptr(self.ctl_sel), sh.first_client, st.n_clients, n, _ptr(sh.dev), rows,
                             _ptr(self._kde_rec), _ptr(self._kde_dens), _ptr(self.mse))
        _lib.check(self.lib.fm_kde_js(C.byref(a), self._stream()), "kde_js")
        self.launches += 2
        if self.exchanges_scores:
            return self._publish(self.mse, n, 1)
        return self.mse[:n]

    def _merge_nvls(self, st, sh, ctl: RoundControl, spec: MergeSpec, mse, agg_pos) -> torch.Tensor:
        """Merge through the NVSwitch: local weighted partial + multimem.ld_reduce / multimem.st of this rank's slice
        (``csrc/nvls_kernel.cu``); no selected row crosses NVLink."""
        n, P, nv = len(ctl.sel), self.layout.padded, self.nvls
        arrive_mse, expect_mse = self._arrive(1) if mse is not None else (0, 0)
        self._nvls_seq += 1
        self.seq[2] += 1
        a = _lib.FmMergeNvlsArgs(_ptr(st.params), self.comm.world, self.comm.rank, sh.n_global, _ptr(self.ctl_sel), n,
                                 _ptr(mse), 1 if spec.update_type in ("mse_avg", "fusion_avg") else 0,
                                 1 if spec.compat_alias else 0, _ptr(agg_pos), _ptr(st.agg_count), _ptr(self.weights),
                                 arrive_mse, expect_mse, nv.ptr, nv.mc_ptr, nv.mc_ptr + 4 * P, _ptr(self.ctl_slab.table),
                                 self.ctl_arrive_ptr + 4 * _lib.ARRIVE_PARTIAL, self.comm.world * self._nvls_seq,
                                 _ptr(self._nvls_counter))
        _lib.check(self.lib.fm_merge_nvls(C.byref(a), self._stream()), "merge_nvls")
        # theta is complete once every rank has stored its slice: order the stream (the verification forward reads it)
        _lib.check(self.lib.fm_flag_wait(self.ctl_arrive_ptr + 4 * _lib.ARRIVE_DONE, self.comm.world * self.seq[2],
                                         self._stream()), "flag_wait")
        self.launches += 2
        return self.theta

    def merge(self, st, sh, ctl: RoundControl, spec: MergeSpec, mse, agg_pos, rows=None) -> torch.Tensor:
        n = len(ctl.sel)
        if self.nvls is not None:
            return self._merge_nvls(st, sh, ctl, spec, mse, agg_pos)
        a = _lib.FmMergeArgs(_ptr(self.peer_table), self.comm.world, self.comm.rank, sh.n_global,
                             _ptr(self.ctl_sel), n, _ptr(mse),
                             1 if spec.update_type in ("mse_avg", "fusion_avg") else 0,
                             1 if spec.compat_alias else 0, _ptr(agg_pos), _ptr(st.agg_count), _ptr(self.theta),
                             _ptr(self.weights), 0, 0, 0, 0)
        if self.exchanges_scores:
            if mse is not None:
                a.arrive, a.expect = self._arrive(1)
            # after its reduction every rank bumps arrive[2] everywhere; verify waits for all of them before
            # any owner overwrites rows that peers may still be pulling
            a.done_ctl, a.done_counter = _ptr(self.ctl_slab.table), _ptr(self.done_counter)
            self.seq[2] += 1
        _lib.check(self.lib.fm_merge(C.byref(a), self._stream()), "merge")
        self.launches += 1
        return self.theta

    # ------------------------------------------------------------------ T3
    def verify_update(self, st, sh, ctl: RoundControl, spec: MergeSpec, theta, agg_pos) -> None:
        if spec.verification_method == "dev":
            R = self.dev_rows
            tiles = self._tiles(("vdev", R), lambda: make_tiles([(0, 0, R, 0)], self.tile))
            X, off, shared = sh.dev, self._tiles(("voff", R), lambda: torch.tensor([0, R], dtype=torch.int32)), 1
            if self.verify_sse.numel() < R:
                self.verify_sse = torch.zeros(R, dtype=torch.float32, device=self.device)
        elif ctl.verify_client >= 0:
            g = ctl.verify_client
            r0, R = sh.valid.off_host[g], sh.valid.count(g)
            tiles = self._tiles(("vshared", g), lambda: make_tiles([(0, r0, R, 0)], self.tile))
            X, off, shared = sh.valid.x, self._tiles(("voff", R), lambda: torch.tensor([0, R], dtype=torch.int32)), 1
        else:
            tiles, X, off, shared = self.tiles_verify_own, sh.valid.x, self.verify_off_own, 0
        self._score(theta, 1, 0, None, X, tiles, self.verify_sse, None)
        a = _lib.FmVerifyArgs(_ptr(st.params), _ptr(st.anchor), _ptr(st.hist), _ptr(st.hist_perf),
                              _ptr(st.hist_valid), _ptr(st.rejected), st.n_clients, sh.first_client, _ptr(theta),
                              _ptr(agg_pos), _ptr(self.verify_sse), _ptr(off), shared,
                              spec.verification_threshold, spec.performance_threshold, 0, 0, *self._arrive(2))
        _lib.check(self.lib.fm_verify_update(C.byref(a), self._stream()), "verify_update")
        self.launches += 1

    # ------------------------------------------------------------------ T4
    def evaluate(self, st, sh, model_type: str, want_latents: bool = False, metric: str = "AUC"):
        hybrid = model_type == "hybrid"
        if hybrid:
            self._score(st.params, st.n_clients, sh.first_client, None, sh.train.x, self.tiles_train, None,
                        self.train_lat)
            self._score(st.params, st.n_clients, sh.first_client, None, sh.test.x, self.tiles_test, None,
                        self.test_lat)
        else:
            self._score(st.params, st.n_clients, sh.first_client, None, sh.test.x, self.tiles_test,
                        self.test_sse, self.test_lat if want_latents else None)
        want_scores = metric != "AUC"
        a = _lib.FmAucArgs(1 if hybrid else 0, st.n_clients, _ptr(self.test_sse), _ptr(self.test_lat),
                           _ptr(sh.test.y), _ptr(sh.test.off), _ptr(self.train_lat), _ptr(sh.train.off),
                           _ptr(self.scores_out) if want_scores else 0, _ptr(self.auc), sh.test.max_count(),
                           _ptr(self._auc_scratch(st.n_clients, sh.test.max_count())))
        _lib.check(self.lib.fm_auc(C.byref(a), self._stream()), "auc")
        self.launches += 1
        vals = None
        if want_scores:
            from . import oracle
            vals = [oracle.classification_metrics(torch.nan_to_num(self.scores_out[sh.test.off_host[c]:
                                                                                     sh.test.off_host[c + 1]]),
                                                  sh.test.labels(c))[0] for c in range(st.n_clients)]
        lat = None
        if want_latents:
            host = self.test_lat[:, :7].cpu().numpy().astype("float32")
            lat = [host[sh.test.off_host[c]:sh.test.off_host[c + 1]].copy() for c in range(st.n_clients)]
        return self.auc, vals, lat

    def _auc_scratch(self, n_local: int, max_rows: int) -> Optional[torch.Tensor]:
        """Global key scratch of the AUC kernel for test sets beyond its shared-memory capacity (16,384 rows per
        client); None below that."""
        per = int(self.lib.fm_auc_scratch_keys(max_rows))
        if per == 0:
            return None
        need = per * max(1, n_local)
        if getattr(self, "_auc_keys", None) is None or self._auc_keys.numel() < need:
            self._auc_keys = torch.empty(need, dtype=torch.int64, device=self.device)
        return self._auc_keys

    def evaluate_side(self, st, sh, model_type: str) -> None:
        """Evaluation of the round on a side stream, from a snapshot of the parameters taken now (main
        stream): the caller may enqueue the next round's local training right behind this call -- it only
        keeps 62 of 148 SMs busy, the evaluation forwards fill the rest.  Results land in ``self.auc``;
        fetch them with ``readback_async(..., side=True)``."""
        if getattr(self, "_side", None) is None:
            self._side = torch.cuda.Stream(device=self.device)
            self._ev_snap, self._ev_pack = torch.cuda.Event(), torch.cuda.Event()
            self._eval_params = torch.empty(st.params.shape, dtype=st.params.dtype, device=self.device)
        main = torch.cuda.current_stream(self.device)
        self._eval_params.copy_(st.params)
        self._ev_snap.record(main)
        with torch.cuda.stream(self._side):
            self._side.wait_event(self._ev_snap)
            proxy = types.SimpleNamespace(params=self._eval_params, n_clients=st.n_clients)
            self.evaluate(proxy, sh, model_type, False, "AUC")

    # ------------------------------------------------------------------ gossip
    device_topology = True     # full / ring / random-k graphs are generated inside the gossip kernel

    def gossip(self, st, sh, ctl: RoundControl, spec: MergeSpec, neigh, rows=None, topo=None) -> None:
        """One launch per rank: every selected client merges {itself} + its in-neighbours and gates the result.

        ``topo = (kind, k, seed)`` with kind in full / ring / random-k: the kernel generates the graph itself
        (nothing is built or uploaded on the host); otherwise ``neigh`` (list of neighbour lists) is uploaded as a
        CSR table once per distinct object.  Cross-GPU ordering uses the snapshot-parity protocol of
        ``csrc/gossip_kernel.cu`` (peer-mapped arrival counters), no NCCL call."""
        from ..parallel import topology
        from ..parallel.symmetric import SymmetricSlab
        if spec.update_type == "fusion_avg":
            raise NotImplementedError("fusion_avg weights are an elected-mode merge rule on the CUDA backend "
                                      "(use backend='oracle' for gossip + fusion_avg)")
        comm = self.comm
        if self.snap_slab is None:
            rows_max = self.param_slab.nbytes // (4 * self.layout.padded)
            with torch.cuda.device(self.device):
                self.snap_slab = SymmetricSlab(2 * rows_max, self.layout.padded, comm, rows_max=2 * rows_max)
            self._slabs.append(self.snap_slab)
            self._snap_stride = rows_max * self.layout.padded
            self._gossip_seq = 0
            self._gossip_done = torch.zeros(1, dtype=torch.int32, device=self.device)
        kind, k, seed, off_p, idx_p = 0, 0, 0, 0, 0
        if topo is not None and _lib.TOPO_KIND.get(topo[0], 0) != 0:
            kind, k, seed = _lib.TOPO_KIND[topo[0]], int(topo[1]), int(topo[2])
        else:
            if getattr(self, "_nbr_src", None) is not neigh:     # keep a reference: identity, not id()
                off, idx = topology.to_csr(neigh, include_self=True)
                self._nbr = (off.to(self.device), idx.to(self.device))
                self._nbr_src = neigh
            off_p, idx_p = _ptr(self._nbr[0]), _ptr(self._nbr[1])
        parity = self._gossip_seq & 1
        self._gossip_seq += 1
        snap = self.snap_slab.ptr + 4 * parity * self._snap_stride
        _lib.check(self.lib.fm_gossip_publish(_ptr(st.params), snap, st.n_clients * self.layout.padded,
                                              _ptr(self.ctl_slab.table) if comm.is_dist else 0, comm.world,
                                              _ptr(self._gossip_done), self._stream()), "gossip_publish")
        arrive, expect = (self.ctl_arrive_ptr + 4 * 3, comm.world * self._gossip_seq) if comm.is_dist else (0, 0)
        n = self.n_sel_local
        if n or comm.is_dist:
            a = _lib.FmGossipArgs(_ptr(self.snap_slab.table), comm.world, comm.rank, sh.n_global,
                                  _ptr(st.params), _ptr(st.anchor), _ptr(st.hist), _ptr(st.hist_perf),
                                  _ptr(st.hist_valid), _ptr(st.rejected), st.n_clients, sh.first_client,
                                  _ptr(self.ctl_sel_local), n, off_p, idx_p,
                                  _ptr(sh.valid.x), _ptr(sh.valid.off),
                                  1 if spec.update_type == "mse_avg" else 0, spec.verification_threshold,
                                  spec.performance_threshold, kind, k, ctl.round, seed & (2**64 - 1), parity,
                                  self._snap_stride, arrive, expect)
            _lib.check(self.lib.fm_gossip(C.byref(a), self._stream()), "gossip")
        self.launches += 2

    # ------------------------------------------------------------------ one read-back per round
    def readback(self, st, agg_pos: Optional[torch.Tensor], with_auc: bool):
        """Gather the round's results with one D2H copy and one wait.  Multi-GPU: the per-rank blocks
        are all-gathered on the device first (one small NCCL call), so the lists returned cover ALL clients:
        (aggregator position, epochs_run[N], rejected[N], auc[N] or None)."""
        return self.readback_wait(self.readback_async(st, agg_pos, with_auc))

    def readback_async(self, st, agg_pos: Optional[torch.Tensor], with_auc: bool, side: bool = False):
        """Enqueue the packing of the round's results and their D2H copy; returns a handle for
        ``readback_wait``.  More work (the next round's training) may be enqueued in between: the wait is on
        an event recorded right behind the copy, not on the stream.  ``side``: the AUCs come from
        ``evaluate_side``; what the next training overwrites (epochs, verdicts) is packed on the main stream
        now, the rest on the side stream behind the evaluation."""
        C_, W = st.n_clients, self.comm.world
        if getattr(self, "_rb_dev", None) is None:
            cmax = C_
            if self.comm.is_dist:
                cmax = int(max(self.comm.all_gather_object(C_)))
            self._rb_cmax = cmax
            self._rb_counts = self.comm.all_gather_object(C_) if self.comm.is_dist else [C_]
            self._rb_dev = torch.zeros(2 + 4 * cmax, dtype=torch.int32, device=self.device)
            self._rb_all = torch.zeros(W * (2 + 4 * cmax), dtype=torch.int32, device=self.device)
            self._rb_pin = torch.zeros_like(self._rb_all, device="cpu").pin_memory()
            self._rb_event = torch.cuda.Event()
        cm, rb = self._rb_cmax, self._rb_dev
        rb[0:2].copy_(agg_pos if agg_pos is not None else self.agg_pos.new_full((2,), -1))
        rb[2:2 + C_].copy_(st.epochs_run)
        rb[2 + cm:2 + cm + C_].copy_(st.rejected)
        main = torch.cuda.current_stream(self.device)
        stream = main
        if side:
            self._ev_pack.record(main)
            self._side.wait_event(self._ev_pack)
            stream = self._side
        with torch.cuda.stream(stream):
            if with_auc:
                rb[2 + 2 * cm:2 + 2 * cm + 2 * C_].view(torch.float64).copy_(self.auc[:C_])
            if self.comm.is_dist:
                import torch.distributed as dist
                dist.all_gather_into_tensor(self._rb_all, rb, group=self.comm.group)
                src = self._rb_all
            else:
                src = rb
            self._rb_pin[:src.numel()].copy_(src, non_blocking=True)
            self._rb_event.record(stream)
        return (with_auc, src.numel())

    def readback_wait(self, handle):
        with_auc, _ = handle
        self._rb_event.synchronize()
        cm = self._rb_cmax
        host = self._rb_pin.numpy()
        stride = 2 + 4 * cm
        epochs, rejected, auc = [], [], ([] if with_auc else None)
        for r, cnt in enumerate(self._rb_counts):
            blk = host[r * stride:(r + 1) * stride]
            epochs += blk[2:2 + cnt].tolist()
            rejected += blk[2 + cm:2 + cm + cnt].tolist()
            if with_auc:
                auc += blk[2 + 2 * cm:2 + 2 * cm + 2 * cnt].view(np.float64).tolist()
        return int(host[0]), epochs, rejected, auc

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)

    def close(self) -> None:
        """Release the peer-mapped slabs (raw cudaMalloc + IPC mappings: torch's allocator cannot reclaim them).
        Collective on multi-GPU runs (each slab barriers before it is freed).  The state tensors that alias the
        parameter slab must not be used afterwards."""
        torch.cuda.synchronize(self.device)
        for slab in self._slabs:
            slab.close()
        self._slabs.clear()
        self.param_slab = self.snap_slab = None
