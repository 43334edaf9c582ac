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
