import numpy as np
import pytest

from fedmse_decentralized_b200.utils import (jensenshannon, js_divergence, kde_self_log_density, kl_divergence,
                                             similarity_score)


def test_kde_matches_sklearn():
    from sklearn.neighbors import KernelDensity
    rng = np.random.default_rng(0)
    X = rng.normal(size=(150, 7))
    want = KernelDensity(kernel="gaussian", bandwidth="scott").fit(X).score_samples(X)
    got = kde_self_log_density(X, "scott")
    assert np.allclose(got, want, atol=1e-8)


def test_similarity_score_matches_reference_formula():
    from scipy.spatial.distance import jensenshannon as sp_js
    from sklearn.neighbors import KernelDensity
    rng = np.random.default_rng(1)
    A, B = rng.normal(size=(80, 5)), rng.normal(size=(80, 5)) + 0.5
    dev_scores = KernelDensity(kernel="gaussian", bandwidth="scott").fit(A).score_samples(A)
    want = sp_js(np.exp(dev_scores), np.exp(KernelDensity(kernel="gaussian", bandwidth="scott").fit(B).score_samples(B)))
    assert abs(similarity_score(dev_scores, B) - want) < 1e-10
    assert abs(jensenshannon([0.2, 0.8], [0.2, 0.8])) < 1e-12


def test_gaussian_divergences():
    m0, m1 = np.zeros(3), np.ones(3)
    c0, c1 = np.eye(3), 2 * np.eye(3)
    assert abs(kl_divergence(m0, c0, m0, c0)) < 1e-12
    want = 0.5 * (3 / 2 + 3 / 2 - 3 + np.log(8))
    assert abs(kl_divergence(m0, c0, m1, c1) - want) < 1e-12
    assert js_divergence(m0, c0, m1, c1) > 0 and abs(js_divergence(m0, c0, m0, c0)) < 1e-12


def test_similarity_against_reference_module(reference_modules):
    import importlib
    RU = importlib.import_module("Utils")
    rng = np.random.default_rng(2)
    A = rng.normal(size=(60, 4))
    s = kde_self_log_density(A)
    B = rng.normal(size=(60, 4)) * 1.7
    assert abs(RU.similarity_score(s, B) - similarity_score(s, B)) < 1e-10


def test_fusion_distances_match_scipy_sklearn():
    """fusion_avg weight function on tensors == the reference formula (sklearn KDE + scipy JS distance)
    applied to each model's reconstruction of the dev rows (src/Utils/utils.py:10-24)."""
    import torch
    from scipy.spatial.distance import jensenshannon
    from sklearn.neighbors import KernelDensity
    from fedmse_decentralized_b200.layout import ParamLayout
    from fedmse_decentralized_b200.models import init_packed_models
    from fedmse_decentralized_b200.ops import oracle
    from fedmse_decentralized_b200.utils.similarity import fusion_distances
    torch.manual_seed(0)
    lay = ParamLayout(12, 6, 3)              # low dimension: exp(log-density) stays representable for scipy
    rows = init_packed_models(3, "hybrid", lay, 5.0)
    dev = torch.randn(40, 128)
    d = fusion_distances(rows, dev, lay, max_rows=40)
    X = dev[:, :12].numpy().astype(np.float64)
    ref = KernelDensity(kernel="gaussian", bandwidth="scott").fit(X).score_samples(X)
    for j in range(3):
        rec = oracle.forward(rows[j], dev[:, :12], lay)[1].numpy().astype(np.float64)
        s = KernelDensity(kernel="gaussian", bandwidth="scott").fit(rec).score_samples(rec)
        want = jensenshannon(np.exp(ref), np.exp(s))
        assert abs(float(d[j]) - want) < 1e-5 * max(1.0, want), (j, float(d[j]), want)


def test_random_k_hash_matches_the_device_function():
    """parallel/topology.py restates the gossip kernel's neighbour hash bit for bit (host side of the C ABI)."""
    from fedmse_decentralized_b200.ops import _lib
    from fedmse_decentralized_b200.parallel import topology
    if not _lib.is_available():
        pytest.skip("library not built")
    lib = _lib.load()
    for seed, rnd, g, j, att, n in [(0, 0, 0, 0, 0, 2), (11, 3, 5, 2, 1, 1000), (2**63 + 5, 17, 999, 63, 4, 1000),
                                    (7, 1, 3, 0, 0, 6)]:
        assert topology.random_neighbour(seed, rnd, g, j, att, n) == lib.fm_random_neighbour(seed, rnd, g, j, att, n)
    nb = topology.random_k(50, 4, seed=1, round_idx=2)
    assert all(len(set(r)) == 4 and i not in r for i, r in enumerate(nb))
    assert nb != topology.random_k(50, 4, seed=1, round_idx=3)


def test_numa_binding_is_best_effort():
    """No GPU / no NVML here: the per-rank CPU binding must decline quietly and leave the affinity alone."""
    import os
    from fedmse_decentralized_b200.parallel.world import bind_to_gpu_numa
    before = os.sched_getaffinity(0) if hasattr(os, "sched_getaffinity") else None
    assert bind_to_gpu_numa(0) is None or isinstance(bind_to_gpu_numa(0), int)
    os.environ["FEDMSE_NO_NUMA_BIND"] = "1"
    try:
        assert bind_to_gpu_numa(0) is None
    finally:
        del os.environ["FEDMSE_NO_NUMA_BIND"]
    if before is not None:
        assert os.sched_getaffinity(0) <= before
