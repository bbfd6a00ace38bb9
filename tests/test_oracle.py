"""fp64 oracle learners / preprocessors: they are the truth the kernels are tested against."""
import numpy as np
import pytest

from omldm_b200.oracle.learners import (KMeans, MultiClassPA, NN, ORR, HoeffdingTree, PAClassifier, PARegressor,
                                        SVM, make_oracle_learner, pa_tau)
from omldm_b200.oracle.pipeline import HoldoutSampler
from omldm_b200.oracle.preprocessors import MinMaxScaler, PolynomialFeatures, StandardScaler, poly_factor_table

RNG = np.random.default_rng(0)
W = RNG.standard_normal(20)
X = RNG.standard_normal((3000, 20))


def test_pa_tau_variants():
    assert pa_tau(0.0, 1.0, 0, 1.0) == 0.0
    assert pa_tau(2.0, 4.0, 0, 1.0) == 0.5
    assert pa_tau(2.0, 4.0, 1, 0.1) == 0.1
    assert abs(pa_tau(2.0, 4.0, 2, 1.0) - 2.0 / 4.5) < 1e-12


def test_linear_learners_learn():
    y = np.sign(X @ W)
    for cls, h in ((PAClassifier, {"C": 0.1}), (SVM, {"lambda": 0.001})):
        m = cls(h)
        for i in range(2500):
            m.fit(X[i], y[i])
        assert m.score(X[2500:], y[2500:]) > 0.9
    yr = X @ W + 0.5
    r = PARegressor({"C": 1.0, "epsilon": 0.01})
    for i in range(2500):
        r.fit(X[i], yr[i])
    assert r.score(X[2500:], yr[2500:]) < 0.1


def test_orr_is_exact_ridge_and_additive():
    yr = X @ W + 0.5
    o = ORR({"lambda": 1.0})
    o.fit_batch(X[:1000], yr[:1000])
    Z = np.hstack([X[:1000], np.ones((1000, 1))])
    w_ref = np.linalg.solve(Z.T @ Z + np.eye(21), Z.T @ yr[:1000])
    assert np.allclose(o.w, w_ref, atol=1e-9)
    a, b = ORR({"lambda": 1.0}), ORR({"lambda": 1.0})
    a.fit_batch(X[:500], yr[:500]); b.fit_batch(X[500:1000], yr[500:1000])
    merged = ORR({"lambda": 1.0}); merged.ensure_init(20)
    merged.set_flat(a.get_flat() + b.get_flat())                                    # replicas merge by addition
    assert np.allclose(merged.w, o.w, atol=1e-9)
    p = o.to_parameters(); o2 = ORR({"lambda": 1.0}); o2.from_parameters(p)
    assert np.allclose(o2.w, o.w, atol=1e-8)


def test_multiclass_kmeans_nn_ht():
    Wm = RNG.standard_normal((4, 20)); ym = np.argmax(X @ Wm.T, 1).astype(float)
    m = MultiClassPA({"nClasses": 4, "C": 1.0})
    for i in range(2500):
        m.fit(X[i], ym[i])
    assert m.score(X[2500:], ym[2500:]) > 0.8
    cent = RNG.standard_normal((5, 8)) * 6; lab = RNG.integers(0, 5, 2000)
    Xk = cent[lab] + RNG.standard_normal((2000, 8))
    k = KMeans({"k": 5})
    for x in Xk:
        k.fit(x, 0.0)
    assert abs(k.S[:, -1].sum() - 2000) < 1e-9 and k.score(Xk) < 60.0
    k2 = KMeans({"k": 5}); k2.from_parameters(k.to_parameters())
    assert np.allclose(k2.centroids(), k.centroids())
    n = NN({"hiddenLayers": [16], "learningRate": 0.05}); yr = X @ W
    first = n.fit_batch(X[:64], yr[:64]) / 64
    for _ in range(5):
        for i in range(0, 2500, 64):
            n.fit_batch(X[i:i + 64], yr[i:i + 64])
    assert n.score(X[2500:], yr[2500:]) ** 2 < first
    flat = n.get_flat().copy(); n.set_flat(flat); assert np.allclose(n.get_flat(), flat)
    ht = HoeffdingTree({"nClasses": 2, "gracePeriod": 100}); yh = (X[:, 0] > 0.3).astype(float)
    for i in range(2500):
        ht.fit(X[i], yh[i])
    assert ht.score(X[2500:], yh[2500:]) > 0.95 and len(ht.feature) >= 3
    assert set(ht.data_structure()) >= {"feature", "threshold", "left", "right", "leaves"}


def test_preprocessors():
    assert poly_factor_table(2, 2).tolist() == [[0, 0], [0, 1], [1, 1], [0, -1], [1, -1]]
    assert PolynomialFeatures({"degree": 2}).transform(np.array([2.0, 3.0])).tolist() == [4.0, 6.0, 9.0, 2.0, 3.0]
    assert PolynomialFeatures({"degree": 2}).out_dim(28) == 434
    a, b = StandardScaler(), StandardScaler()
    for x in X[:500]:
        a.fit(x)
    b.fit_batch(X[:200]); b.fit_batch(X[200:500])
    assert np.allclose(a.mean, b.mean) and np.allclose(a.m2, b.m2)                 # Welford == Chan merge
    z = a.transform_batch(X[:500]); assert abs(z.mean()) < 1e-9 and abs(z.std(ddof=1) - 1) < 5e-2
    mm = MinMaxScaler(); zz = mm.fit_transform_batch(X[:100]); assert zz.min() == 0.0 and zz.max() == 1.0


def test_holdout_sampler_rule():
    h = HoldoutSampler(4)
    trained = []
    for c in range(40):
        r = h.route(np.array([float(c)]), 0.0)
        if r is not None:
            trained.append(int(r[0][0]))
    # positions 8,9 of every 10 are held out; once 4 are held, each new one evicts (trains) the oldest
    assert trained[:8] == list(range(8)) and 8 not in trained[:10]
    assert [int(p[0][0]) for p in h.test_set] == [28, 29, 38, 39]
    assert 8 in trained and 9 in trained and 18 in trained and 19 in trained


def test_factory_rejects_unknown():
    with pytest.raises(KeyError):
        make_oracle_learner("XGBoost")
