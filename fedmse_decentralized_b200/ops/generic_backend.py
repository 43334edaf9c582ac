"""Backend for autoencoders of any width: generic sm_100a kernels for the heavy ops, torch for the small ones.

The specialised kernels behind ``CudaBackend`` are laid out for the reference's default 115-27-7 model and batches
of at most 32 rows.  The reference's model classes take ``hidden_neus`` / ``latent_dim``
(``src/Model/Shrink_Autoencoder.py:125-127``) and the thesis trains with batch 64; this backend serves those
settings on the GPU: local training (``fm_train_generic``) and the scoring forwards of the MSE-weight merge and
of the evaluation (``fm_score_generic``) are shape-generic kernels of ``csrc/generic_kernels.cu``; the small
steps (vote on one validation matrix, weight normalisation, weighted sum, verification decision, rank
statistic) reuse the fp32 torch restatement of ``OracleBackend`` on the device tensors.  Multi-GPU runs go through
the engine's gather path (``Comm.gather_rows``), like the oracle backend.
"""
from __future__ import annotations

import ctypes as C
from typing import List, Optional, Sequence

import torch

from ..layout import D_PAD, ParamLayout
from . import _lib, oracle
from .backend import MergeSpec, OracleBackend, RoundControl
from .oracle import TrainHP

GTILE = 16


def _ptr(t: Optional[torch.Tensor]) -> int:
    return 0 if t is None else t.data_ptr()


def generic_train(lib, layout: ParamLayout, st, sh, sel_local_dev: torch.Tensor, n: int, hp: TrainHP, stream: int,
                  scratch_holder) -> None:
    """One launch of ``fm_train_generic`` for ``n`` selected local clients (ids in ``sel_local_dev``)."""
    per = int(lib.fm_train_generic_scratch_floats(layout.dim, layout.hidden, layout.latent, hp.batch_size))
    buf = getattr(scratch_holder, "_gen_scratch", None)
    if buf is None or buf.numel() < per * n:
        buf = torch.empty(per * max(n, 1), dtype=torch.float32, device=st.params.device)
        scratch_holder._gen_scratch = buf
    a = _lib.FmTrainGenericArgs(
        _ptr(st.params), _ptr(st.adam_m), _ptr(st.adam_v), _ptr(st.adam_t), _ptr(st.anchor), _ptr(st.best),
        _ptr(st.loss_log), _ptr(st.epochs_run), _ptr(st.trained_rounds),
        _ptr(sh.train.x), _ptr(sh.train.off), _ptr(sh.valid.x), _ptr(sh.valid.off),
        _ptr(sel_local_dev), n, sh.first_client, hp.epochs, hp.batch_size, hp.patience, st.loss_log.shape[1],
        hp.lr, hp.shrink_lambda, hp.fedprox_mu, layout.dim, layout.hidden, layout.latent, _ptr(buf), per)
    _lib.check(lib.fm_train_generic(C.byref(a), stream), "train_generic")


class GenericCudaBackend(OracleBackend):
    name = "cuda-generic"

    def __init__(self, layout: ParamLayout, device):
        super().__init__(layout, device)
        if self.device.type != "cuda":
            raise RuntimeError("GenericCudaBackend needs a CUDA device")
        if layout.dim > D_PAD:
            raise ValueError(f"dim_features {layout.dim} exceeds the shard pitch {D_PAD}")
        self.lib = _lib.load()
        self.launches = 0

    def _stream(self) -> int:
        return torch.cuda.current_stream(self.device).cuda_stream

    # ------------------------------------------------------------------ T1
    def train(self, st, sh, sel_local: Sequence[int], hp: TrainHP) -> None:
        st.epochs_run.zero_()
        if not len(sel_local):
            return
        sel = torch.tensor(list(sel_local), dtype=torch.int32, device=self.device)
        generic_train(self.lib, self.layout, st, sh, sel, len(sel_local), hp, self._stream(), self)
        self.launches += 1

    # ------------------------------------------------------------------ scoring forward of many (model, row) pairs
    def _score(self, slab: torch.Tensor, segments, X: torch.Tensor, want_sse: bool, want_lat: bool):
        """``segments`` = [(local row of slab, row0, nrows, out0)]; returns (sse[total] or None, lat[total, Z] or None)."""
        L = self.layout
        total = sum(s[2] for s in segments)
        tiles = []
        for c, row0, nrows, out0 in segments:
            for s0 in range(0, nrows, GTILE):
                tiles.append((c, row0 + s0, min(GTILE, nrows - s0), out0 + s0))
        if not tiles:
            return (torch.zeros(0, device=self.device) if want_sse else None,
                    torch.zeros(0, L.latent, device=self.device) if want_lat else None)
        t = torch.tensor(tiles, dtype=torch.int32, device=self.device).reshape(-1, 4)
        sse = torch.empty(total, dtype=torch.float32, device=self.device) if want_sse else None
        lat = torch.empty(total, L.latent, dtype=torch.float32, device=self.device) if want_lat else None
        a = _lib.FmScoreGenericArgs(_ptr(slab), 0, 0, slab.shape[0], _ptr(X), _ptr(t), t.shape[0], _ptr(sse), _ptr(lat),
                                    L.latent, L.dim, L.hidden, L.latent)
        _lib.check(self.lib.fm_score_generic(C.byref(a), self._stream()), "score_generic")
        self.launches += 1
        return sse, lat

    # ------------------------------------------------------------------ T2: MSE weights (the O(models x rows) part)
    def merge_scores(self, st, sh, ctl: RoundControl) -> torch.Tensor:
        out = torch.zeros(len(ctl.sel), dtype=torch.float32, device=st.device)
        R = sh.dev.shape[0]
        own = [(pos, g - sh.first_client) for pos, g in enumerate(ctl.sel) if self._owned(sh, g)]
        if own:
            sse, _ = self._score(st.params, [(c, 0, R, k * R) for k, (_, c) in enumerate(own)], sh.dev, True, False)
            mse = sse.view(len(own), R).double().sum(1) / (R * self.layout.dim)
            out[torch.tensor([pos for pos, _ in own], device=st.device)] = mse.float()
        return out

    # ------------------------------------------------------------------ T4
    def evaluate(self, st, sh, model_type: str, want_latents: bool = False, metric: str = "AUC"):
        L = self.layout
        C_ = st.n_clients
        hybrid = model_type == "hybrid"
        seg_te = [(c, sh.test.off_host[c], sh.test.count(c), sh.test.off_host[c]) for c in range(C_)]
        sse_te, lat_te = self._score(st.params, seg_te, sh.test.x, not hybrid, hybrid or want_latents)
        lat_tr = None
        if hybrid:
            seg_tr = [(c, sh.train.off_host[c], sh.train.count(c), sh.train.off_host[c]) for c in range(C_)]
            _, lat_tr = self._score(st.params, seg_tr, sh.train.x, False, True)
        aucs = torch.zeros(C_, dtype=torch.float32, device=st.device)
        aucs64: List[float] = []
        latents = [] if want_latents else None
        for c in range(C_):
            lo, hi = sh.test.off_host[c], sh.test.off_host[c + 1]
            if hybrid:
                ztr = lat_tr[sh.train.off_host[c]:sh.train.off_host[c + 1]]
                mu = ztr.mean(dim=0)
                sd = ztr.std(dim=0, unbiased=False)
                sd = torch.where(sd == 0, torch.ones_like(sd), sd)
                score = torch.linalg.vector_norm((lat_te[lo:hi] - mu) / sd, dim=1)
            else:
                score = sse_te[lo:hi] / L.dim
            if metric == "classification":
                val = oracle.classification_metrics(torch.nan_to_num(score), sh.test.labels(c))[0]
            else:
                val = oracle.auc_from_scores(score, sh.test.labels(c))
            aucs64.append(val)
            aucs[c] = val
            if want_latents:
                latents.append(lat_te[lo:hi].detach().cpu().numpy().astype("float32"))
        return aucs, aucs64, latents

    def synchronize(self) -> None:
        torch.cuda.synchronize(self.device)
