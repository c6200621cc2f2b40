"""Tests against the committed fixtures in tests/golden/ (see tests/golden/make_golden.py for what they are).

  upstream_doctest_kats.json      values transcribed from upstream pyspark doctests (the only upstream-derived numbers
                                  here; soft: upstream prints ~3 digits)
  oracle_regression_vectors.npz   outputs of our own CPU oracle on seeded inputs -- a regression pin and a committed
                                  target for the CUDA path, NOT evidence of parity with Spark

CPU tests hold the oracle to both.  GPU tests compare the CUDA path, through the C ABI and the public API, with the
committed vectors and do not import the oracle at all.  Tolerances: labels / counts exact; everything floating point
1e-5 relative (north_star), the upstream printed digits to their printed precision.
"""
import json
import os

import numpy as np
import pytest

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def _kats():
    with open(os.path.join(GOLD, "upstream_doctest_kats.json")) as f:
        return json.load(f)


def _vec():
    return np.load(os.path.join(GOLD, "oracle_regression_vectors.npz"))


# ------------------------------------------------------------------ CPU: the oracle against the fixtures
def test_oracle_meets_the_upstream_doctest_values():
    import oracle

    K = _kats()
    c = K["kmeans_four_points"]
    X = np.array(c["points"], dtype=np.float64)
    run = oracle.kmeans_lloyd(X, X[[0, 2]].copy(), max_iter=10, tol=1e-4)
    got = run["centers"][np.argsort(run["centers"][:, 0])]
    np.testing.assert_allclose(got, c["centers"], atol=1e-12)
    assert run["cost_hist"][-1] == pytest.approx(c["cost"], abs=1e-12)
    c = K["lr_weighted"]
    fit = oracle.lr_fit(np.array(c["X"], dtype=np.float32), np.array(c["y"]), np.array(c["w"], dtype=np.float64),
                        reg_param=c["regParam"], tol=1e-12)
    np.testing.assert_allclose(fit["coefficients"], c["upstream_printed"]["coef"], atol=1e-3)   # what upstream prints
    assert fit["intercept"] == pytest.approx(c["upstream_printed"]["intercept"], abs=1e-3)
    np.testing.assert_allclose(fit["coefficients"], c["coef"], atol=1e-6)                       # reproduced digits
    assert fit["intercept"] == pytest.approx(c["intercept"], abs=1e-6)
    assert fit["objectiveHistory"][-1] == pytest.approx(c["objective"], abs=1e-7)
    c = K["lr_two_points_maxiter5"]
    fit = oracle.lr_fit(np.array(c["X"], dtype=np.float32), np.array(c["y"]), None, max_iter=c["maxIter"],
                        reg_param=c["regParam"])
    assert fit["coefficients"][0] == pytest.approx(c["upstream_printed"]["coef"][0], abs=0.05)
    assert fit["intercept"] == pytest.approx(c["upstream_printed"]["intercept"], abs=0.01)
    assert fit["coefficients"][0] == pytest.approx(c["coef"][0], rel=1e-8)
    assert fit["intercept"] == pytest.approx(c["intercept"], rel=1e-8)
    np.testing.assert_allclose(fit["objectiveHistory"], c["objective_history"], atol=1e-6)


def test_oracle_reproduces_its_committed_regression_vectors():
    import oracle

    z = _vec()
    lab, cost, sums, counts = oracle.kmeans_step(z["km_X"].astype(np.float64), z["km_C0"])
    assert np.array_equal(lab, z["km_labels"]) and np.array_equal(counts, z["km_counts"])
    assert cost == pytest.approx(float(z["km_cost"]), rel=1e-13)
    np.testing.assert_allclose(sums, z["km_sums"], rtol=1e-13)
    run = oracle.kmeans_lloyd(z["km_X"].astype(np.float64), z["km_C0"], max_iter=4, tol=0.0)
    np.testing.assert_allclose(run["centers"], z["km_centers_after4"], rtol=1e-12)
    np.testing.assert_allclose(run["cost_hist"], z["km_cost_hist"], rtol=1e-12)
    lab2, cost2, _, counts2 = oracle.kmeans_step(z["tc_X"].astype(np.float64), z["tc_C0"])
    assert np.array_equal(lab2, z["tc_labels"]) and np.array_equal(counts2, z["tc_counts"])
    d = z["lr_X"].shape[1]
    loss, g, gb = oracle.lr_binary_lossgrad(z["lr_X"], z["lr_y"], None, z["lr_beta"], float(z["lr_icpt"]), np.ones(d))
    assert loss == pytest.approx(float(z["lr_loss"]), rel=1e-13)
    np.testing.assert_allclose(g, z["lr_grad"], rtol=1e-11, atol=1e-15)
    fit = oracle.lr_fit(z["lr_X"], z["lr_y"].astype(np.int64), None, max_iter=8, reg_param=0.01, tol=1e-6)
    np.testing.assert_allclose(fit["objectiveHistory"], z["lr_fit_hist"], rtol=1e-12)
    np.testing.assert_allclose(fit["coefficients"], z["lr_fit_coef"], rtol=1e-9, atol=1e-12)
    lm, gm, gbm = oracle.lr_multinomial_lossgrad(z["lr_X"], z["mn_y"], None, z["mn_B"], z["mn_b0"], np.ones(d))
    assert lm == pytest.approx(float(z["mn_loss"]), rel=1e-13)
    np.testing.assert_allclose(gm, z["mn_grad"], rtol=1e-11, atol=1e-15)


# ------------------------------------------------------------------ GPU: the CUDA path against the fixtures
def _gpu():
    import torch

    from b200fit import _native as N
    import gpu_util as G

    return torch, N, G


@pytest.mark.gpu
@pytest.mark.parametrize("prefix,algo_name", [("km", "KM_GENERIC"), ("tc", "KM_TC3"), ("tc", "KM_TCB"), ("tc", "KM_TCH")])
def test_cuda_lloyd_step_matches_the_committed_vectors(prefix, algo_name):
    torch, N, G = _gpu()
    z = _vec()
    Xh, C = z[prefix + "_X"], z[prefix + "_C0"]
    n, d = Xh.shape
    X, _, ld = G.upload_matrix(Xh)
    labels, sums, counts, cost, nrows = G.kmeans_step_raw(X, n, d, ld, C, getattr(N, algo_name))
    assert nrows == n
    assert np.array_equal(labels, z[prefix + "_labels"])
    assert np.array_equal(counts, z[prefix + "_counts"])
    assert cost == pytest.approx(float(z[prefix + "_cost"]), rel=1e-5)
    if prefix == "km":
        np.testing.assert_allclose(sums, z["km_sums"], rtol=1e-5, atol=1e-5)


@pytest.mark.gpu
def test_cuda_kmeans_fit_matches_the_committed_run():
    torch, N, G = _gpu()
    import b200fit
    from b200fit.frame import ColumnarFrame

    z = _vec()
    m = b200fit.clustering.KMeans(k=8, maxIter=4, tol=0.0, initialModel=z["km_C0"]).fit(
        ColumnarFrame({"features": z["km_X"]}))
    sc = np.abs(z["km_centers_after4"]).max()
    np.testing.assert_allclose(np.array(m.clusterCenters()), z["km_centers_after4"], rtol=1e-5, atol=1e-5 * sc)
    np.testing.assert_allclose(m.summary.costHistory, z["km_cost_hist"], rtol=1e-5)


@pytest.mark.gpu
def test_cuda_lr_passes_and_fit_match_the_committed_vectors():
    torch, N, G = _gpu()
    import b200fit
    from b200fit.frame import ColumnarFrame

    z = _vec()
    Xh = z["lr_X"]
    n, d = Xh.shape
    X, _, ld = G.upload_matrix(Xh)
    y = torch.from_numpy(z["lr_y"]).to(G.dev())
    loss, g, gb, W = G.lr_lossgrad_raw(X, y, None, n, d, ld, 2, z["lr_beta"], float(z["lr_icpt"]))
    assert W == n and loss == pytest.approx(float(z["lr_loss"]), rel=1e-5)
    np.testing.assert_allclose(g[0], z["lr_grad"], rtol=1e-5, atol=1e-5 * np.abs(z["lr_grad"]).max())
    assert gb[0] == pytest.approx(float(z["lr_grad_icpt"]), rel=1e-5, abs=1e-6)
    ym = torch.from_numpy(z["mn_y"]).to(G.dev())
    lm, gm, gbm, _ = G.lr_lossgrad_raw(X, ym, None, n, d, ld, 5, z["mn_B"], z["mn_b0"])
    assert lm == pytest.approx(float(z["mn_loss"]), rel=1e-5)
    np.testing.assert_allclose(gm, z["mn_grad"], rtol=2e-5, atol=2e-5 * np.abs(z["mn_grad"]).max())
    np.testing.assert_allclose(gbm, z["mn_grad_icpt"], rtol=2e-5, atol=2e-5)
    m = b200fit.classification.LogisticRegression(maxIter=8, regParam=0.01, tol=1e-6).fit(
        ColumnarFrame({"features": Xh, "label": z["lr_y"].astype(np.float64)}))
    h = m.summary.objectiveHistory
    k = min(len(h), len(z["lr_fit_hist"]))
    np.testing.assert_allclose(h[:k], z["lr_fit_hist"][:k], rtol=1e-5)
    sc = np.abs(z["lr_fit_coef"]).max()
    np.testing.assert_allclose(m.coefficients, z["lr_fit_coef"], rtol=1e-4, atol=1e-4 * sc)
    assert m.intercept == pytest.approx(float(z["lr_fit_icpt"]), rel=1e-4, abs=1e-5)


@pytest.mark.gpu
def test_cuda_path_meets_the_upstream_doctest_values():
    torch, N, G = _gpu()
    import b200fit
    from b200fit.frame import ColumnarFrame

    K = _kats()
    c = K["kmeans_four_points"]
    Xh = np.array(c["points"], dtype=np.float32)
    m = b200fit.clustering.KMeans(k=2, initialModel=Xh[[0, 2]].astype(np.float64)).fit(ColumnarFrame({"features": Xh}))
    got = np.array(m.clusterCenters())
    np.testing.assert_allclose(got[np.argsort(got[:, 0])], c["centers"], atol=1e-6)
    assert m.summary.trainingCost == pytest.approx(c["cost"], abs=1e-5)
    c = K["lr_weighted"]
    m = b200fit.classification.LogisticRegression(regParam=c["regParam"], weightCol="wt", tol=1e-12).fit(
        ColumnarFrame({"features": np.array(c["X"], dtype=np.float32), "label": np.array(c["y"], dtype=np.float64),
                       "wt": np.array(c["w"], dtype=np.float32)}))
    np.testing.assert_allclose(m.coefficients, c["upstream_printed"]["coef"], atol=1e-3)   # what upstream prints
    assert m.intercept == pytest.approx(c["upstream_printed"]["intercept"], abs=1e-3)
