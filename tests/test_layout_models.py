import math

import numpy as np
import pytest
import torch

from fedmse_decentralized_b200.layout import DEFAULT_LAYOUT as L, STATE_KEYS, ParamLayout
from fedmse_decentralized_b200.models import (Autoencoder, CentroidBasedOneClassClassifier, Shrink_Autoencoder,
                                              build_model, init_packed_models)
from fedmse_decentralized_b200.ops import oracle


def test_layout_sizes_and_keys():
    assert L.n_params == 6764 and L.padded == 6784 and L.nbytes == 27056
    m = Shrink_Autoencoder(115, 115)
    sd = m.state_dict()
    assert tuple(sd.keys()) == STATE_KEYS
    assert [tuple(v.shape) for v in sd.values()] == L.shapes


def test_pack_unpack_roundtrip():
    m = Shrink_Autoencoder(115, 115)
    row = m.pack()
    assert row.shape == (6784,) and float(row[6764:].abs().sum()) == 0
    sd = L.unpack(row)
    for k, v in m.state_dict().items():
        assert torch.equal(sd[k], v)
    m2 = Autoencoder(115, 115)
    m2.load_packed(row)
    x = torch.randn(5, 115)
    assert torch.allclose(m2(x)[1], m(x)[1])


def test_forward_and_losses_match_oracle():
    torch.manual_seed(0)
    x = torch.randn(12, 115)
    for mt, lam in (("hybrid", 5.0), ("autoencoder", 0.0)):
        m = build_model(mt, shrink_lambda=5)
        z, xh, loss = m(x)
        z2, xh2 = oracle.forward(m.pack(), x, L)
        assert torch.allclose(z, z2, atol=1e-6) and torch.allclose(xh, xh2, atol=1e-6)
        assert torch.allclose(loss, oracle.batch_loss(m.pack(), x, L, lam), atol=1e-6)
    # un-squared shrink norm (quirk Q15)
    m = Shrink_Autoencoder(115, 115, shrink_lambda=3)
    z, xh, loss = m(x)
    want = ((x - xh) ** 2).mean() + 3 * z.norm(dim=1).mean()
    assert torch.allclose(loss, want, atol=1e-6)


def test_init_distribution_and_determinism():
    torch.manual_seed(3)
    a = init_packed_models(3, "hybrid", L)
    torch.manual_seed(3)
    b = init_packed_models(3, "hybrid", L)
    assert torch.equal(a, b) and not torch.equal(a[0], a[1])
    W1, b1, W2, b2, W3, b3, W4, b4 = L.views(a[0])
    for W in (W1, W2, W3, W4):
        assert float(W.abs().max()) <= 1 / math.sqrt(W.shape[1]) + 1e-7
    for bb in (b1, b2, b3, b4):
        assert float(bb.abs().sum()) == 0


def test_init_matches_reference_rng_order(reference_modules):
    RM = reference_modules["Model"]
    torch.manual_seed(11)
    ref = [RM.Shrink_Autoencoder(input_dim=115, output_dim=115, shrink_lambda=5) for _ in range(2)]
    torch.manual_seed(11)
    ours = init_packed_models(2, "hybrid", L, 5)
    for c in range(2):
        for k, v in ref[c].state_dict().items():
            assert torch.equal(L.unpack(ours[c])[k], v), k


def test_centroid_matches_sklearn_pipeline():
    from sklearn.preprocessing import StandardScaler
    rng = np.random.default_rng(0)
    tr = rng.normal(size=(200, 7)); tr[:, 3] = 2.5      # constant column -> scale 1
    te = rng.normal(size=(50, 7)) * 3
    cen = CentroidBasedOneClassClassifier().fit(tr)
    sk = StandardScaler().fit(tr)
    want = np.linalg.norm(sk.transform(te), axis=1)
    assert np.allclose(cen.get_density(te), want)
    assert cen.predict(te).dtype == bool


def test_other_dims_supported_by_oracle():
    lay = ParamLayout(46, 27, 7)       # cic-2023 has 46 features (src/main.py:62 comment)
    m = build_model("hybrid", 46)
    assert m.pack().shape[0] == lay.padded
