"""Plain-PyTorch fp32 restatement of every fused op (the numerics oracle).

Each function here computes, with stock torch ops on whatever device the
tensors live on, exactly what one of the sm_100a kernels in ``csrc/`` computes.
It is (a) the reference the GPU tests compare the kernels against, (b) the
backend that keeps the whole framework runnable on a CPU-only box, and (c) a
line-by-line restatement of the reference's math, which
``tests/test_reference_parity.py`` checks against the reference's own modules.

Reference sites: local training ``src/Trainer/client_trainer.py:360-419``; losses
``src/Model/Shrink_Autoencoder.py:152-153``, ``src/Model/AutoEncoder.py:148``;
vote score ``client_trainer.py:208-247``; merges ``:107-134``; verification
``src/Trainer/model_verifier.py:29-99``; evaluation
``src/Evaluator/evaluator.py:50-119`` and ``src/Model/Centroid.py:15-35``.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import List, Optional, Sequence, Tuple

import torch

from ..layout import ParamLayout

ADAM_B1, ADAM_B2, ADAM_EPS = 0.9, 0.999, 1e-8


@dataclass
class TrainHP:
    epochs: int = 5
    batch_size: int = 12
    lr: float = 1e-3
    shrink_lambda: float = 5.0      # 0 for the plain AE
    fedprox_mu: float = 0.0         # >0 only for update_type == "fedprox"
    patience: int = 1


# --------------------------------------------------------------------------- forward
def forward(row: torch.Tensor, x: torch.Tensor, layout: ParamLayout):
    """(latent, reconstruction) of rows ``x[:, :dim]`` under the packed model ``row``."""
    W1, b1, W2, b2, W3, b3, W4, b4 = layout.views(row)
    h1 = torch.relu(x @ W1.t() + b1)
    z = h1 @ W2.t() + b2
    h2 = torch.relu(z @ W3.t() + b3)
    return z, h2 @ W4.t() + b4


def batch_loss(row, x, layout: ParamLayout, lam: float, mu: float = 0.0,
               anchor: Optional[torch.Tensor] = None):
    z, xh = forward(row, x, layout)
    loss = torch.mean((x - xh) ** 2)
    if lam != 0.0:
        loss = loss + lam * (torch.linalg.vector_norm(z, dim=1).sum() / x.shape[0])
    if mu != 0.0:
        n = layout.n_params
        loss = loss + mu * torch.sum((row[:n] - anchor[:n]) ** 2)
    return loss


# --------------------------------------------------------------------------- T1
def train_client(params: torch.Tensor, m: torch.Tensor, v: torch.Tensor, t: int,
                 anchor: torch.Tensor, x_train: torch.Tensor, x_valid: torch.Tensor,
                 layout: ParamLayout, hp: TrainHP):
    """One ``ClientTrainer.run`` call. Updates ``params/m/v`` in place.

    Returns ``(t, best_row, [(train_loss, valid_loss), ...])``.  Mirrors the
    reference exactly: fixed batch order, partial last batch, Adam with torch's
    operation order, FedProx term in training *and* validation loss, best-epoch
    snapshot on strict improvement, stop once ``patience`` non-improving epochs
    accumulated, in-memory weights are those of the last epoch (quirk Q12).
    """
    D, n = layout.dim, layout.n_params
    B = hp.batch_size
    xt, xv = x_train[:, :D], x_valid[:, :D]
    best = params.clone()
    min_valid, worse = float("inf"), 0
    log: List[Tuple[float, float]] = []
    w = params  # updated in place
    for _ in range(hp.epochs):
        tot, nb = 0.0, 0
        for s in range(0, xt.shape[0], B):
            leaf = w.detach().clone().requires_grad_(True)
            loss = batch_loss(leaf, xt[s:s + B], layout, hp.shrink_lambda, hp.fedprox_mu, anchor)
            (g,) = torch.autograd.grad(loss, leaf)
            g[n:] = 0
            t += 1
            with torch.no_grad():
                m.lerp_(g, 1 - ADAM_B1)
                v.mul_(ADAM_B2).addcmul_(g, g, value=1 - ADAM_B2)
                bc1 = 1 - ADAM_B1 ** t
                bc2 = 1 - ADAM_B2 ** t
                denom = (v.sqrt() / math.sqrt(bc2)).add_(ADAM_EPS)
                w.addcdiv_(m, denom, value=-(hp.lr / bc1))
            tot += float(loss.detach())
            nb += 1
        train_loss = tot / max(nb, 1)
        vt, vb = 0.0, 0
        with torch.no_grad():
            for s in range(0, xv.shape[0], B):
                vt += float(batch_loss(w, xv[s:s + B], layout, hp.shrink_lambda, hp.fedprox_mu, anchor))
                vb += 1
        valid_loss = vt / vb if vb else float("nan")
        log.append((train_loss, valid_loss))
        if vb == 0:
            continue          # run(train_loader, valid_loader=None): all epochs, no early stop, nothing saved
        if valid_loss < min_valid:
            min_valid, worse = valid_loss, 0
            best.copy_(w)
        else:
            worse += 1
            if worse >= hp.patience:
                break
    return t, best, log


# --------------------------------------------------------------------------- T2 pieces
def vote_normalise(V: torch.Tensor, dim: int) -> torch.Tensor:
    """Column-normalise the vote matrix: (V - mean) / (unbiased std + 1e-8) (quirk Q4)."""
    out = torch.zeros_like(V)
    X = V[:, :dim]
    out[:, :dim] = (X - X.mean(dim=0, keepdim=True)) / (X.std(dim=0, keepdim=True) + 1e-8)
    return out


def row_sse(row: torch.Tensor, x: torch.Tensor, layout: ParamLayout) -> torch.Tensor:
    """Per-row sum of squared reconstruction errors."""
    xx = x[:, :layout.dim]
    _, xh = forward(row, xx, layout)
    return ((xx - xh) ** 2).sum(dim=1)


def chunked_mse(row: torch.Tensor, x: torch.Tensor, layout: ParamLayout, chunk: int) -> torch.Tensor:
    """Mean over consecutive ``chunk``-row blocks of each block's element-mean squared error.

    ``chunk >= rows`` gives the plain full-batch ``MSELoss(reduction='mean')``.
    """
    sse = row_sse(row, x, layout)
    R = sse.shape[0]
    if chunk <= 0 or chunk >= R:
        return sse.sum() / (R * layout.dim)
    means = [sse[s:s + chunk].sum() / (min(chunk, R - s) * layout.dim) for s in range(0, R, chunk)]
    return torch.stack(means).mean()


def mse_scores(slab: torch.Tensor, idx: Sequence[int], x: torch.Tensor, layout: ParamLayout,
               chunk: int = 0) -> torch.Tensor:
    return torch.stack([chunked_mse(slab[int(j)], x, layout, chunk) for j in idx]) \
        if len(idx) else torch.zeros(0, device=slab.device)


def merge_weights(update_type: str, mse: Optional[torch.Tensor], n: int, agg_pos: int = -1,
                  compat_alias: bool = False, device=None) -> torch.Tensor:
    """Normalised merge weights over the ``n`` selected models (aggregator included).

    ``avg`` / ``fedprox``: uniform.  ``mse_avg``: ``s_j = 1/MSE_j`` normalised.
    ``compat_alias`` reproduces quirk Q6 of the reference's ``fed_mse_avg``: the
    aggregator (position ``agg_pos``) loads every candidate into its own live
    tensors, so its own slot is scored with the model loaded just before its turn
    and finally contributes a duplicate of the last-loaded peer.  Expressed on the
    source models this moves the aggregator's weight onto that peer.
    """
    if update_type in ("avg", "fedprox"):
        return torch.full((n,), 1.0 / n, dtype=torch.float32, device=device)
    s = 1.0 / mse.to(torch.float32)
    if compat_alias and agg_pos >= 0 and n > 1:
        s = s.clone()
        if agg_pos > 0:
            s[agg_pos] = s[agg_pos - 1]           # scored while holding model a-1
        last = n - 1 if agg_pos != n - 1 else n - 2
        w = s / s.sum()
        moved = w[agg_pos].clone()
        w[agg_pos] = 0
        w[last] = w[last] + moved
        return w
    return s / s.sum()


def weighted_reduce(slab: torch.Tensor, idx: Sequence[int], w: torch.Tensor) -> torch.Tensor:
    """theta = sum_j w_j * slab[idx_j] accumulated in selection order (fp32)."""
    out = torch.zeros(slab.shape[1], dtype=torch.float32, device=slab.device)
    for k, j in enumerate(idx):
        out += slab[int(j)] * w[k]
    return out


# --------------------------------------------------------------------------- T3
def param_change(old: torch.Tensor, new: torch.Tensor, layout: ParamLayout) -> torch.Tensor:
    """Sum over the 8 tensors of the per-tensor L2 norm of the difference."""
    d = old - new
    return sum(torch.linalg.vector_norm(seg.reshape(-1)) for seg in layout.views(d))


def performance(row: torch.Tensor, x: torch.Tensor, layout: ParamLayout) -> torch.Tensor:
    return 1.0 / (1.0 + chunked_mse(row, x, layout, 0))


# --------------------------------------------------------------------------- T4
def auc_from_scores(score: torch.Tensor, label: torch.Tensor) -> float:
    """Tie-aware ROC-AUC = Mann-Whitney U / (n+ n-), computed exactly in integers.

    Equals sklearn ``auc(*roc_curve(y, s)[:2])`` (trapezoid over tied thresholds).
    Non-finite scores are mapped like ``np.nan_to_num`` first
    (``src/Evaluator/evaluator.py:22-24``).
    """
    s = torch.nan_to_num(score.detach().to(torch.float32).cpu())
    y = label.detach().cpu() > 0.5
    n_pos, n_neg = int(y.sum()), int((~y).sum())
    if n_pos == 0 or n_neg == 0:
        return float("nan")
    order = torch.argsort(s, stable=True)
    ss, yy = s[order], y[order].to(torch.int64)
    # group ties
    new_grp = torch.ones_like(ss, dtype=torch.bool)
    new_grp[1:] = ss[1:] != ss[:-1]
    gid = torch.cumsum(new_grp.to(torch.int64), 0) - 1
    ng = int(gid[-1]) + 1
    pos_g = torch.zeros(ng, dtype=torch.int64).index_add_(0, gid, yy)
    neg_g = torch.zeros(ng, dtype=torch.int64).index_add_(0, gid, 1 - yy)
    neg_below = torch.cumsum(neg_g, 0) - neg_g
    u2 = (pos_g * (2 * neg_below + neg_g)).sum()       # 2*U, exact
    return float(u2.item()) / (2.0 * n_pos * n_neg)


def anomaly_scores(row: torch.Tensor, x_test: torch.Tensor, x_train: Optional[torch.Tensor],
                   layout: ParamLayout, model_type: str):
    """Per-row anomaly score and test latents.

    ``autoencoder``: mean squared reconstruction error per row.  ``hybrid``:
    Euclidean norm of the test latent standardised by the train latents'
    mean / population std (zero std -> 1).
    """
    D = layout.dim
    z_te, xh = forward(row, x_test[:, :D], layout)
    if model_type == "autoencoder":
        return ((x_test[:, :D] - xh) ** 2).mean(dim=1), z_te
    z_tr, _ = forward(row, x_train[:, :D], layout)
    mu = z_tr.mean(dim=0)
    sd = z_tr.std(dim=0, unbiased=False)
    sd = torch.where(sd == 0, torch.ones_like(sd), sd)
    return torch.linalg.vector_norm((z_te - mu) / sd, dim=1), z_te


def classification_metrics(score: torch.Tensor, label: torch.Tensor, threshold: float = 0.5):
    """F1 / precision / recall of ``score > threshold`` (``src/Evaluator/evaluator.py:30-47``)."""
    pred = score.detach().cpu() > threshold
    y = label.detach().cpu() > 0.5
    tp = int((pred & y).sum()); fp = int((pred & ~y).sum()); fn = int((~pred & y).sum())
    precision = tp / (tp + fp) if tp + fp else 0.0
    recall = tp / (tp + fn) if tp + fn else 0.0
    f1 = 2 * precision * recall / (precision + recall) if precision + recall else 0.0
    return f1, precision, recall
