"""Similarity utilities (the reference's orphaned ``Utils`` package).

``similarity_score`` is the weight function of the historical ``fusion_avg``
merge rule whose source was deleted from the reference (last
rows): fit a Gaussian KDE with Scott's bandwidth on a generated data set, take
its log-density at its own points, and return the Jensen-Shannon *distance*
between ``exp`` of those scores and the dev-set scores
(``src/Utils/utils.py:10-24``).  ``kl_divergence`` / ``js_divergence`` are the closed
forms for two Gaussians (``:26-52``).

Implemented directly with numpy/torch (pairwise squared distances by one GEMM +
log-sum-exp), so it runs on the GPU for large sets.
"""
from __future__ import annotations

import math

import numpy as np
import torch


def scott_bandwidth(n: int, d: int) -> float:
    """sklearn's ``bandwidth='scott'``: ``n ** (-1 / (d + 4))``."""
    return float(n) ** (-1.0 / (d + 4))


def kde_self_log_density(X, bandwidth="scott", device=None, block: int = 4096) -> np.ndarray:
    """log p_KDE(x_i) for every row of ``X`` under the Gaussian KDE fitted on ``X`` itself."""
    Xt = torch.as_tensor(np.asarray(X), dtype=torch.float64, device=device)
    n, d = Xt.shape
    h = scott_bandwidth(n, d) if bandwidth == "scott" else float(bandwidth)
    log_norm = -0.5 * d * math.log(2 * math.pi) - d * math.log(h) - math.log(n)
    sq = (Xt * Xt).sum(1)
    out = torch.empty(n, dtype=torch.float64, device=Xt.device)
    for s in range(0, n, block):
        xb = Xt[s:s + block]
        d2 = (sq[s:s + block, None] + sq[None, :] - 2.0 * xb @ Xt.t()).clamp_min_(0)
        out[s:s + block] = torch.logsumexp(-0.5 * d2 / (h * h), dim=1) + log_norm
    return out.cpu().numpy()


def jensenshannon(p, q) -> float:
    """scipy's ``jensenshannon``: sqrt of the JS divergence (natural log) of the normalised vectors."""
    p = np.asarray(p, dtype=np.float64)
    q = np.asarray(q, dtype=np.float64)
    p = p / p.sum()
    q = q / q.sum()
    m = 0.5 * (p + q)

    def _rel(a, b):
        mask = a > 0
        return np.sum(a[mask] * np.log(a[mask] / b[mask]))

    return float(np.sqrt(max(0.0, 0.5 * _rel(p, m) + 0.5 * _rel(q, m))))


def similarity_score(dev_kde_scores, dataset_2, device=None) -> float:
    kde2 = kde_self_log_density(dataset_2, "scott", device)
    return jensenshannon(np.exp(np.asarray(dev_kde_scores)), np.exp(kde2))


def kl_divergence(p_mean, p_cov, q_mean, q_cov) -> float:
    k = p_mean.shape[0]
    q_inv = np.linalg.inv(q_cov)
    diff = q_mean - p_mean
    return float(0.5 * (np.trace(q_inv @ p_cov) + diff.T @ q_inv @ diff - k
                        + np.log(np.linalg.det(q_cov) / np.linalg.det(p_cov))))


def js_divergence(p_mean, p_cov, q_mean, q_cov) -> float:
    m_mean = 0.5 * (p_mean + q_mean)
    m_cov = 0.5 * (p_cov + q_cov)
    return 0.5 * (kl_divergence(p_mean, p_cov, m_mean, m_cov) + kl_divergence(q_mean, q_cov, m_mean, m_cov))


def fusion_weights(dev_kde_scores, generated_sets, device=None) -> np.ndarray:
    """``fusion_avg`` weights: inverse JS distance of each client's generated data to the dev set."""
    d = np.array([similarity_score(dev_kde_scores, g, device) for g in generated_sets])
    w = 1.0 / (d + 1e-12)
    return w / w.sum()


# --------------------------------------------------------------------------- fusion_avg as a live merge rule
def _js_distance_from_logs(a: torch.Tensor, b: torch.Tensor) -> torch.Tensor:
    """``jensenshannon(exp(a), exp(b))`` for log-score vectors, batched over leading dims of ``b``.

    scipy normalises both vectors to sum 1 first, so only ``softmax`` of the log-scores matters; computing it
    as a softmax keeps 115-dimensional KDE log-densities (hundreds of nats below zero) from underflowing."""
    p = torch.softmax(a, dim=-1)
    q = torch.softmax(b, dim=-1)
    m = 0.5 * (p + q)

    def rel(x, y):
        t = torch.where(x > 0, x * (torch.log(x.clamp_min(1e-300)) - torch.log(y.clamp_min(1e-300))),
                        torch.zeros_like(x))
        return t.sum(-1)

    return torch.sqrt((0.5 * rel(p, m) + 0.5 * rel(q, m)).clamp_min(0))


def _kde_self_log_density_t(X: torch.Tensor) -> torch.Tensor:
    """Batched ``kde_self_log_density`` on tensors ``[..., n, d]`` (float64, Scott bandwidth)."""
    n, d = X.shape[-2], X.shape[-1]
    h = scott_bandwidth(n, d)
    log_norm = -0.5 * d * math.log(2 * math.pi) - d * math.log(h) - math.log(n)
    sq = (X * X).sum(-1)
    d2 = (sq[..., :, None] + sq[..., None, :] - 2.0 * X @ X.transpose(-1, -2)).clamp_min_(0)
    return torch.logsumexp(-0.5 * d2 / (h * h), dim=-1) + log_norm


def fusion_distances(rows: torch.Tensor, dev: torch.Tensor, layout, max_rows: int = 512,
                     model_block: int = 16) -> torch.Tensor:
    """Per-model ``similarity_score`` for the ``fusion_avg`` merge rule, on the tensors' device.

    The source of ``GlobalAggregator.fusion_avg`` was deleted from the reference; what survives is its weight
    function (``src/Utils/utils.py:10-24``) and the names in the leftover bytecode.  Here the
    "generated data set" of client j is the dev set as *reconstructed by client j's autoencoder* (same number
    of points as the dev set, which ``jensenshannon`` needs), and the score is the JS distance between the KDE
    self-densities of the dev rows and of that reconstruction.  ``max_rows`` caps the O(n^2 d) pairwise work by
    using the first rows of the (already shuffled) dev matrix.  Returns ``[len(rows)]`` float32 distances; the
    merge weight is ``(1/d_j) / sum_k (1/d_k)`` (``fusion_weights``)."""
    from ..ops import oracle
    D = layout.dim
    X = dev[:max_rows, :D].to(torch.float32)
    ref = _kde_self_log_density_t(X.to(torch.float64))
    out = torch.empty(rows.shape[0], dtype=torch.float32, device=rows.device)
    for s in range(0, rows.shape[0], model_block):
        rec = torch.stack([oracle.forward(r, X, layout)[1] for r in rows[s:s + model_block]])
        dens = _kde_self_log_density_t(rec.to(torch.float64))
        out[s:s + model_block] = _js_distance_from_logs(ref, dens).to(torch.float32) + 1e-12
    return out
