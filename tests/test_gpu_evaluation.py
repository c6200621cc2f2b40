"""GPU parity tests for the evaluators against oracle/evaluate.py, which
tests/test_oracle_kat.py pins against scikit-learn.  Tolerances: the silhouette kernel works in fp32 per row
(fp64 sums) -> 1e-5 absolute; the ranking / confusion-matrix metrics are fp64 -> 1e-12."""
import numpy as np
import pytest
import torch

import fake_widget as fw
import oracle
from oracle import synth

pytestmark = pytest.mark.gpu
SEED = 20240611


def _mods():
    import b200fit
    from b200fit import _native as N
    from b200fit.frame import ColumnarFrame

    return b200fit, N, ColumnarFrame


@pytest.mark.parametrize("n,d,k", [(2, 1, 2), (500, 3, 4), (4099, 16, 8), (3000, 130, 5), (20000, 64, 256)])
@pytest.mark.parametrize("weighted", [False, True])
def test_silhouette_matches_oracle(n, d, k, weighted):
    b200fit, N, ColumnarFrame = _mods()
    rng = np.random.default_rng(n + d + k)
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    lab = rng.integers(0, k, size=n)
    lab[0], lab[-1] = 0, k - 1
    if n > 10:
        lab[lab == 1] = 0   # an empty cluster
        lab[7] = 1          # ... now a singleton
    wh = rng.uniform(0.5, 2.0, size=n).astype(np.float32) if weighted else None
    cols = {"features": Xh, "prediction": lab.astype(np.int32)}
    kw = {}
    if weighted:
        cols["wt"] = wh
        kw["weightCol"] = "wt"
    got = b200fit.evaluation.ClusteringEvaluator(**kw).evaluate(ColumnarFrame(cols))
    ref = oracle.silhouette_sqeuclid(Xh.astype(np.float64), lab, wh)
    assert got == pytest.approx(ref, abs=1e-5)


def test_silhouette_through_the_c_abi_two_shards_accumulate():
    """The statistics buffer is accumulated: two shards through b2f_silhouette_stats equal one pass (what the
    all-reduce relies on), and the scores of the shards add up."""
    b200fit, N, _ = _mods()
    import gpu_util as G

    L = N.lib()
    n, d, k = 3001, 20, 6
    Xh = synth.mixture(n, 0, d, SEED, k, 2.0)
    lab = np.random.default_rng(1).integers(0, k, size=n).astype(np.int32)
    X, _, ld = G.upload_matrix(Xh)
    labd = torch.from_numpy(lab).to(G.dev())
    stats = torch.zeros(k * (d + 2), dtype=torch.float64, device=G.dev())
    h = n // 3
    for a, b in ((0, h), (h, n)):
        N.check(L.b2f_silhouette_stats(N.ptr(X[a:]), b - a, d, ld, N.ptr(labd[a:]), None, k, N.ptr(stats), None,
                                       G.stream()), "stats")
    S = stats.cpu().numpy().reshape(k, d + 2)
    for c in range(k):
        m = lab == c
        np.testing.assert_allclose(S[c, :d], Xh[m].astype(np.float64).sum(0), rtol=1e-12, atol=1e-9)
        assert S[c, d] == pytest.approx((Xh[m].astype(np.float64) ** 2).sum(), rel=1e-12)
        assert S[c, d + 1] == m.sum()
    scratch = torch.empty(int(L.b2f_silhouette_scratch_bytes(k, d)), dtype=torch.uint8, device=G.dev())
    out = torch.zeros(2, dtype=torch.float64, device=G.dev())
    for a, b in ((0, h), (h, n)):
        N.check(L.b2f_silhouette_score(N.ptr(X[a:]), b - a, d, ld, N.ptr(labd[a:]), None, k, N.ptr(stats),
                                       N.ptr(scratch), N.ptr(out), G.stream()), "score")
    tot, cnt = out.cpu().numpy()
    assert cnt == n and tot / cnt == pytest.approx(oracle.silhouette_sqeuclid(Xh.astype(np.float64), lab), abs=1e-5)
    bad = torch.zeros(1, dtype=torch.int32, device=G.dev())
    labd[5] = k
    N.check(L.b2f_silhouette_stats(N.ptr(X), n, d, ld, N.ptr(labd), None, k, N.ptr(stats), N.ptr(bad), G.stream()), "s")
    assert int(bad.item()) == 1


def test_evaluation_widget_path_after_kmeans_and_logreg():
    """Estimator.fit -> Model.transform -> Evaluation widget 'Apply' (one call per metric name parsed out of the
    metricName doc, reference: orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:42-55)."""
    b200fit, N, ColumnarFrame = _mods()
    n, d, k = 6000, 16, 8
    Xh = synth.mixture(n, 0, d, SEED, k, 3.0)
    df = ColumnarFrame({"features": Xh})
    model = b200fit.clustering.KMeans(k=k, seed=3, maxIter=10).fit(df)
    out = model.transform(df)
    vals = fw.press_apply_evaluator(b200fit.evaluation.ClusteringEvaluator, out)
    lab = out.col("prediction").cpu().numpy()
    assert set(vals) == {"silhouette"}
    assert vals["silhouette"] == pytest.approx(oracle.silhouette_sqeuclid(Xh.astype(np.float64), lab), abs=1e-5)

    rng = np.random.default_rng(4)
    Xl = rng.normal(size=(5000, 10)).astype(np.float32)
    wt = rng.normal(size=10)
    yl = (rng.uniform(size=5000) < 1 / (1 + np.exp(-(Xl @ wt)))).astype(np.float64)
    dfl = ColumnarFrame({"features": Xl, "label": yl})
    m = b200fit.classification.LogisticRegression(maxIter=20, regParam=0.01).fit(dfl)
    pl = m.transform(dfl)
    raw = pl.col("rawPrediction").cpu().numpy().astype(np.float64)
    vals = fw.press_apply_evaluator(b200fit.evaluation.BinaryClassificationEvaluator, pl)
    assert vals["areaUnderROC"] == pytest.approx(oracle.area_under_roc(raw[:, 1], yl), rel=1e-12)
    assert vals["areaUnderPR"] == pytest.approx(oracle.area_under_pr(raw[:, 1], yl), rel=1e-12)
    assert vals["areaUnderROC"] > 0.8
    vals = fw.press_apply_evaluator(b200fit.evaluation.MulticlassClassificationEvaluator, pl)
    ref = oracle.multiclass_metrics(pl.col("prediction").cpu().numpy(), yl)
    for name in ("f1", "accuracy", "weightedPrecision", "weightedRecall"):
        assert vals[name] == pytest.approx(ref[name], rel=1e-12)


def test_binary_evaluator_ties_weights_and_errors():
    b200fit, N, ColumnarFrame = _mods()
    rng = np.random.default_rng(9)
    y = (rng.uniform(size=3000) < 0.3).astype(np.float64)
    sc = np.round(y * 0.7 + rng.normal(size=3000), 1)  # heavy ties
    w = rng.uniform(0.5, 2.0, size=3000).astype(np.float32)
    df = ColumnarFrame({"rawPrediction": sc, "label": y, "wt": w})
    ev = b200fit.evaluation.BinaryClassificationEvaluator(weightCol="wt")
    assert ev.evaluate(df) == pytest.approx(oracle.area_under_roc(sc, y, w.astype(np.float64)), rel=1e-9)
    assert ev.evaluate(df, {ev.metricName: "areaUnderPR"}) == pytest.approx(
        oracle.area_under_pr(sc, y, w.astype(np.float64)), rel=1e-9)
    with pytest.raises(ValueError):
        b200fit.evaluation.BinaryClassificationEvaluator().evaluate(
            ColumnarFrame({"rawPrediction": sc, "label": np.ones(3000)}))
    with pytest.raises(ValueError):
        b200fit.evaluation.ClusteringEvaluator(metricName="nope").evaluate(
            ColumnarFrame({"features": np.zeros((4, 2), np.float32), "prediction": np.array([0, 1, 0, 1])}))
