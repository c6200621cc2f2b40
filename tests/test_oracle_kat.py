"""Pins for the oracle: the three upstream pyspark doctest known-answer values
(BASELINE.md section 5 / SURVEY.md Appendix B) plus scikit-learn cross-checks at
the optimum.  These are SOFT pins (upstream prints 3 digits); real-Spark parity
stays unpinned -- see oracle/__init__.py."""
import numpy as np
import pytest

import oracle
from oracle.kmeans import find_closest_scalar


def test_kmeans_four_point_kat():
    X = np.array([[0.0, 0.0], [1.0, 1.0], [9.0, 8.0], [8.0, 9.0]])
    C0 = X[[0, 2]]
    r = oracle.kmeans_lloyd(X, C0, max_iter=20, tol=1e-4)
    got = r["centers"][np.argsort(r["centers"][:, 0])]
    np.testing.assert_allclose(got, [[0.5, 0.5], [8.5, 8.5]], atol=1e-12)
    assert r["trainingCost"] == pytest.approx(2.0, abs=1e-12)
    assert r["converged"]


def test_lr_weighted_kat():
    # pyspark >= 2.1 doctest: DenseVector([-1.080..., -0.646...]), intercept 3.112...
    X = np.array([[0.0, 5.0], [1.0, 2.0], [2.0, 1.0], [3.0, 3.0]])
    y = np.array([1, 0, 1, 0])
    w = np.array([1.0, 2.0, 3.0, 4.0])
    r = oracle.lr_fit(X, y, w, reg_param=0.01, max_iter=100, tol=1e-12)
    np.testing.assert_allclose(r["coefficients"], [-1.08072651, -0.64629055], atol=2e-6)
    assert r["intercept"] == pytest.approx(3.11276685, abs=2e-6)
    assert r["objectiveHistory"][-1] == pytest.approx(0.52728645, abs=1e-7)


def test_lr_trajectory_kat():
    # pyspark <= 2.0 doctest LogisticRegression(maxIter=5, regParam=0.01): 5.5..., -2.68...
    X = np.array([[1.0], [0.0]])
    y = np.array([1, 0])
    r = oracle.lr_fit(X, y, None, reg_param=0.01, max_iter=5)
    assert r["coefficients"][0] == pytest.approx(5.538218804, abs=1e-8)
    assert r["intercept"] == pytest.approx(-2.685823155, abs=1e-8)
    np.testing.assert_allclose(
        r["objectiveHistory"], [0.693147, 0.460384, 0.271403, 0.155430, 0.140877, 0.137702], atol=1e-6
    )
    assert r["totalIterations"] == 5
    assert r["evals"] == 6  # one data pass per iteration + the initial one


def test_vectorised_assign_equals_scalar_loop():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(200, 5))
    C = rng.normal(size=(7, 5))
    C[3] = C[1]  # exact tie -> lowest index wins
    X[10] = C[1]
    lab, dist = oracle.kmeans_assign(X, C)
    xn = np.sqrt((X * X).sum(1))
    cn = np.sqrt((C * C).sum(1))
    for i in range(X.shape[0]):
        j, dd = find_closest_scalar(X[i], xn[i], C, cn)
        assert j == lab[i]
        assert dd == pytest.approx(dist[i], abs=1e-12)
    assert lab[10] == 1


def test_kmeans_empty_cluster_keeps_centre_and_cost_monotone():
    rng = np.random.default_rng(1)
    X = rng.normal(size=(500, 3))
    C0 = np.vstack([X[:3], [[100.0, 100.0, 100.0]]])
    r = oracle.kmeans_lloyd(X, C0, max_iter=10, tol=0.0)
    np.testing.assert_array_equal(r["centers"][3], C0[3])
    c = r["cost_hist"]
    assert all(c[i + 1] <= c[i] * (1 + 1e-12) for i in range(len(c) - 1))
    assert sum(r["counts_hist"][-1]) == 500


def test_lr_binary_matches_sklearn_optimum():
    sk = pytest.importorskip("sklearn.linear_model")
    rng = np.random.default_rng(2)
    n, d = 400, 6
    X = rng.normal(size=(n, d)) * rng.uniform(0.5, 2.0, size=d)
    wt = rng.normal(size=d)
    y = (rng.uniform(size=n) < 1 / (1 + np.exp(-(X @ wt + 0.3)))).astype(int)
    lam = 0.05
    r = oracle.lr_fit(X, y, None, reg_param=lam, max_iter=200, tol=1e-12)
    # Spark penalises standardized coefficients: fit sklearn on X/std, C = 1/(n*lam)
    Xs = X / r["std"]
    m = sk.LogisticRegression(C=1.0 / (n * lam), tol=1e-12, max_iter=5000).fit(Xs, y)
    np.testing.assert_allclose(r["coefficients"] * r["std"], m.coef_[0], atol=2e-5)
    assert r["intercept"] == pytest.approx(m.intercept_[0], abs=2e-5)


def test_lr_multinomial_matches_sklearn_optimum():
    sk = pytest.importorskip("sklearn.linear_model")
    rng = np.random.default_rng(3)
    n, d, K = 600, 5, 4
    X = rng.normal(size=(n, d))
    Wt = rng.normal(size=(K, d))
    y = np.array([rng.choice(K, p=np.exp(z - z.max()) / np.exp(z - z.max()).sum()) for z in X @ Wt.T])
    lam = 0.02
    r = oracle.lr_fit(X, y, None, reg_param=lam, max_iter=300, tol=1e-12, family="multinomial")
    Xs = X / r["std"]
    m = sk.LogisticRegression(C=1.0 / (n * lam), tol=1e-12, max_iter=5000).fit(Xs, y)
    np.testing.assert_allclose(r["coefficients"] * r["std"][None, :], m.coef_, atol=5e-5)
    np.testing.assert_allclose(r["intercept"], m.intercept_ - m.intercept_.mean(), atol=5e-5)


def test_owlqn_elastic_net_optimum_is_stationary():
    """OWL-QN restatement is un-probed upstream; check its fixed point instead: at
    the returned point the pseudo-gradient of (loss + L2 + L1) vanishes."""
    rng = np.random.default_rng(4)
    n, d, K = 500, 8, 3
    X = rng.normal(size=(n, d))
    Wt = rng.normal(size=(K, d)) * (rng.uniform(size=(K, d)) < 0.5)
    y = np.argmax(X @ Wt.T + rng.gumbel(size=(n, K)), axis=1)
    lam, alpha = 0.05, 0.5
    r = oracle.lr_fit(X, y, None, reg_param=lam, elastic_net=alpha, max_iter=500, tol=1e-12,
                      family="multinomial", return_states=True)
    st = r["states"][-1]
    assert np.linalg.norm(st["adj_grad"]) < 1e-5
    assert (r["coefficients"] == 0.0).sum() > 0  # L1 produced exact zeros
    h = r["objectiveHistory"]
    assert h[-1] < h[0]


def test_assemble_and_scaler():
    a = np.array([1, 2, 3], dtype=np.int32)
    b = np.array([True, False, True])
    c = np.array([0.5, np.nan, 2.5], dtype=np.float32)
    # handleInvalid: 'error' (default) rejects NaN and null, 'keep' lets NaN through, Inf is an ordinary value
    with pytest.raises(ValueError):
        oracle.assemble([a, b, c])
    X = oracle.assemble([a, b, c], handle_invalid="keep")
    assert X.dtype == np.float64 and X.shape == (3, 3)
    assert np.isnan(X[1, 2]) and X[2, 0] == 3.0 and X[0, 1] == 1.0
    with pytest.raises(ValueError):
        oracle.assemble([np.array([1.0, None], dtype=object)])
    assert np.isinf(oracle.assemble([np.array([1.0, np.inf, -np.inf])])[1:, 0]).all()
    Z = np.array([[1.0, 5.0], [2.0, 5.0], [4.0, 5.0]])
    mean, std = oracle.standard_scaler_fit(Z)
    np.testing.assert_allclose(std, [np.std(Z[:, 0], ddof=1), 0.0])
    T = oracle.standard_scaler_transform(Z, mean, std)
    np.testing.assert_allclose(T[:, 0], Z[:, 0] / std[0])
    assert (T[:, 1] == 0).all()



def test_evaluator_restatements_against_scikit_learn():
    """Pins oracle/evaluate.py: Spark's statistics form of the squared-Euclidean silhouette equals the pairwise
    definition; exact ROC area and the weighted multiclass metrics equal scikit-learn's."""
    from sklearn.metrics import (accuracy_score, f1_score, precision_score, recall_score, roc_auc_score,
                                 silhouette_score)

    rng = np.random.default_rng(0)
    lab = rng.integers(0, 5, 600)
    X = rng.normal(size=(600, 7)) + lab[:, None] * 1.5
    assert oracle.silhouette_sqeuclid(X, lab) == pytest.approx(silhouette_score(X, lab, metric="sqeuclidean"), rel=1e-12)
    y = (rng.uniform(size=600) < 0.4).astype(float)
    sc = np.round(y * 0.8 + rng.normal(size=600), 1)  # many tied scores
    assert oracle.area_under_roc(sc, y) == pytest.approx(roc_auc_score(y, sc), rel=1e-12)
    w = rng.uniform(0.5, 2.0, 600)
    assert oracle.area_under_roc(sc, y, w) == pytest.approx(roc_auc_score(y, sc, sample_weight=w), rel=1e-12)
    # PR area: hand-checked 4-point case, curve (0,1) (0.5,1) (0.5,0.5) (1,2/3) (1,0.5)
    assert oracle.area_under_pr([0.9, 0.8, 0.7, 0.6], [1, 0, 1, 0]) == pytest.approx(0.5 * 1.0 + 0.5 * (0.5 + 2 / 3) / 2)
    p, t = rng.integers(0, 4, 600), rng.integers(0, 4, 600)
    m = oracle.multiclass_metrics(p, t)
    assert m["accuracy"] == pytest.approx(accuracy_score(t, p))
    assert m["f1"] == pytest.approx(f1_score(t, p, average="weighted"))
    assert m["weightedPrecision"] == pytest.approx(precision_score(t, p, average="weighted"))
    assert m["weightedRecall"] == pytest.approx(recall_score(t, p, average="weighted"))
