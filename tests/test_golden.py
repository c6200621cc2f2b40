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
    np.testing.assert_allclose(gm, z["mn_grad"], rtol=1e-5, atol=1e-5 * np.abs(z["mn_grad"]).max())
    np.testing.assert_allclose(gbm, z["mn_grad_icpt"], rtol=1e-5, atol=1e-5)
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


# ------------------------------------------------------------------ slots for golden vectors from REAL Spark
# tools/spark_oracle.py (run where pyspark + a JVM exist) writes tests/golden/spark_cfg<N>.npz: centres after every
# Lloyd iteration / coefficients after every optimizer iteration on the seeded Philox data.  Until someone commits
# those files the tests below skip, and the status of the oracle stays "parity unpinned" (DESIGN.md section 2).
def _spark_golden(cfg):
    path = os.path.join(GOLD, "spark_cfg%d.npz" % cfg)
    if not os.path.exists(path):
        pytest.skip("no Spark golden committed (run tools/spark_oracle.py where pyspark is installed)")
    return np.load(path, allow_pickle=False)


def _spark_inputs(z, cfg):
    """Rebuild the inputs of a golden file from its recorded seeds with the repository's own generator."""
    from oracle import synth

    n, d, seed = int(z["n"]), int(z["d"]), int(z["data_seed"])
    if cfg in (1, 2, 5):
        col = z["col_scale"] if "col_scale" in z.files else None
        X = synth.mixture(n, 0, d, seed, int(z["k"]), float(z["center_scale"]), col)
        return X, None
    return synth.logistic_features(n, 0, d, seed, z["col_scale"]), z["y"]


def _trajectory_comparable(z):
    major, minor = (int(v) for v in str(z["spark_version"]).split(".")[:2])
    return (major, minor) < (3, 2)  # Spark >= 3.2 centres the features: same optimum, different path


def test_spark_oracle_script_generates_the_same_data_as_the_oracle():
    """tools/spark_oracle.py carries its own copy of the Philox generator (it must run without this repository);
    the copy has to stay bit-identical to oracle/synth.py, which mirrors the device generator."""
    import importlib.util

    from oracle import synth

    spec = importlib.util.spec_from_file_location(
        "spark_oracle", os.path.join(os.path.dirname(GOLD), "..", "tools", "spark_oracle.py"))
    mod = importlib.util.module_from_spec(spec)
    spec.loader.exec_module(mod)  # imports numpy only; pyspark is imported inside main()
    col = np.linspace(0.5, 2.0, 7).astype(np.float32)
    assert np.array_equal(mod.mixture(501, 1000, 7, mod.SEED, 11, 2.0, col), synth.mixture(501, 1000, 7, mod.SEED, 11, 2.0, col))
    assert np.array_equal(mod.logistic_features(300, 5, 9, mod.SEED + 1, None), synth.logistic_features(300, 5, 9, mod.SEED + 1))
    y = mod.logistic_labels(mod.logistic_features(300, 0, 9, 3), 3, np.ones((4, 9)) * 0.1, np.zeros(4), 4)
    assert set(np.unique(y)) <= {0.0, 1.0, 2.0, 3.0}


@pytest.mark.parametrize("cfg", [1, 2, 5])
def test_oracle_kmeans_against_spark_golden(cfg):
    import oracle

    z = _spark_golden(cfg)
    X, _ = _spark_inputs(z, cfg)
    X = X.astype(np.float64)
    if cfg == 5:
        mean, std = oracle.standard_scaler_fit(X)
        np.testing.assert_allclose(std, z["scaler_std"], rtol=1e-12)
        X = oracle.standard_scaler_transform(X, mean, std)
    C0 = X[z["init_rows"]]
    T = len(z["centers_hist"])
    run = oracle.kmeans_lloyd(X, C0, max_iter=T, tol=float(z["tol"]))
    for t in range(min(T, len(run["centers_hist"]))):
        np.testing.assert_allclose(run["centers_hist"][t], z["centers_hist"][t], rtol=1e-9, atol=1e-12)
        assert run["cost_hist"][t] == pytest.approx(float(z["cost_of_assignment"][t]), rel=1e-9)


@pytest.mark.parametrize("cfg", [3, 4])
def test_oracle_logreg_against_spark_golden(cfg):
    import oracle

    z = _spark_golden(cfg)
    X, y = _spark_inputs(z, cfg)
    fit = oracle.lr_fit(X, y.astype(np.int64), None, max_iter=int(z["total_iterations"]), reg_param=float(z["reg_param"]),
                        elastic_net=float(z["elastic_net"]), tol=1e-6, family="binomial" if cfg == 3 else "multinomial")
    sc = np.abs(z["coef_hist"][-1]).max()
    np.testing.assert_allclose(np.atleast_2d(fit["coefficients"]), z["coef_hist"][-1], rtol=1e-5, atol=1e-5 * sc)
    np.testing.assert_allclose(np.atleast_1d(fit["intercept"]), z["icpt_hist"][-1], rtol=1e-5, atol=1e-5)
    if _trajectory_comparable(z):
        h = z["objective_history"]
        assert len(fit["objectiveHistory"]) == len(h)
        np.testing.assert_allclose(fit["objectiveHistory"], h, rtol=1e-9)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [1, 2, 5])
def test_cuda_kmeans_against_spark_golden(cfg):
    torch, N, G = _gpu()
    import b200fit
    from b200fit.frame import ColumnarFrame

    z = _spark_golden(cfg)
    X, _ = _spark_inputs(z, cfg)
    frame = ColumnarFrame({"raw": b200fit.feature.pack_columns([X])})
    if cfg == 5:
        sm = b200fit.feature.StandardScaler(inputCol="raw", outputCol="features").fit(frame)
        np.testing.assert_allclose(sm.std, z["scaler_std"], rtol=1e-9)
        frame = sm.transform(frame)
        C0 = frame.col("features").numpy().astype(np.float64)[z["init_rows"]]
    else:
        frame = ColumnarFrame({"features": frame.col("raw")})
        C0 = X.astype(np.float64)[z["init_rows"]]
    T = len(z["centers_hist"])
    for t in range(1, T + 1):  # as the script does: a run of t iterations gives the centres after iteration t
        m = b200fit.clustering.KMeans(k=int(z["k"]), maxIter=t, tol=float(z["tol"]), initialModel=C0).fit(frame)
        sc = np.abs(z["centers_hist"][t - 1]).max()
        np.testing.assert_allclose(np.stack(m.clusterCenters()), z["centers_hist"][t - 1], rtol=1e-5, atol=1e-5 * sc)
        assert m.summary.trainingCost == pytest.approx(float(z["cost_of_assignment"][t - 1]), rel=1e-5)


@pytest.mark.gpu
@pytest.mark.parametrize("cfg", [3, 4])
def test_cuda_logreg_against_spark_golden(cfg):
    torch, N, G = _gpu()
    import b200fit
    from b200fit.frame import ColumnarFrame

    z = _spark_golden(cfg)
    X, y = _spark_inputs(z, cfg)
    m = b200fit.classification.LogisticRegression(
        maxIter=int(z["total_iterations"]), regParam=float(z["reg_param"]), elasticNetParam=float(z["elastic_net"]),
        tol=1e-6, family="binomial" if cfg == 3 else "multinomial").fit(
        ColumnarFrame({"features": b200fit.feature.pack_columns([X]), "label": y}))
    sc = np.abs(z["coef_hist"][-1]).max()
    np.testing.assert_allclose(m.coefficientMatrix, z["coef_hist"][-1], rtol=1e-5, atol=1e-5 * sc)
    np.testing.assert_allclose(m.interceptVector, z["icpt_hist"][-1], rtol=1e-5, atol=1e-5)
    if _trajectory_comparable(z):
        h = z["objective_history"]
        k = min(len(h), len(m.summary.objectiveHistory))
        np.testing.assert_allclose(m.summary.objectiveHistory[:k], h[:k], rtol=1e-5)
