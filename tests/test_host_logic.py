"""CPU tests of the host side: parameter marshalling through the fake widget, the frame,
the product's own L-BFGS / OWL-QN against the oracle's restatement, C-ABI symbol export."""
import ctypes
import inspect
import os
import re

import numpy as np
import pytest

import oracle
from oracle.optim import LBFGS as OLBFGS
from oracle.optim import OWLQN as OOWLQN

import b200fit
from b200fit import _native
from b200fit.frame import ColumnarFrame
from b200fit.optim import LBFGS, OWLQN

import fake_widget as fw

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def test_cabi_library_exports_every_declared_symbol():
    hdr = open(os.path.join(ROOT, "include", "b200fit.h")).read()
    declared = set(re.findall(r"\b(b2f_[a-z0-9_]+)\s*\(", hdr))
    assert len(declared) >= 20
    lib = ctypes.CDLL(_native.LIB_PATH)  # load only; no compute without a GPU
    for name in declared:
        assert hasattr(lib, name), "missing export %s" % name
    assert declared == set(_native.PROTOTYPES), "ctypes prototypes out of sync with the header"
    lib.b2f_version.restype = ctypes.c_int
    assert lib.b2f_version() >= 100


def test_estimator_discovery_and_description():
    # exact sets: the abstract bases imported into the modules (Estimator, Transformer, Model) must NOT show up as
    # selectable entries -- pyspark's are ABCs and the widget filters with inspect.isabstract
    assert set(fw.list_estimators(b200fit.clustering)) == {"KMeans"}
    assert set(fw.list_estimators(b200fit.classification)) == {"LogisticRegression"}
    assert set(fw.list_estimators(b200fit.feature)) == {"StandardScaler"}
    assert set(fw.list_transformers(b200fit.feature)) == {"VectorAssembler"}
    for base in (b200fit.base.Estimator, b200fit.base.Transformer, b200fit.base.Model, b200fit.evaluation.Evaluator):
        assert inspect.isabstract(base)
        with pytest.raises(TypeError):
            base()
    name, doc, table = fw.describe(b200fit.clustering.KMeans)
    assert name == "b200fit.clustering.KMeans" and doc
    assert list(table)[:3] == ["featuresCol", "predictionCol", "k"]
    assert table["k"][1] == 2 and table["tol"][1] == 1e-4 and table["initMode"][1] == "k-means||"
    assert all(len(v) == 3 and v[2] for v in table.values()), "every ctor kwarg needs a Param doc"
    _, _, t2 = fw.describe(b200fit.classification.LogisticRegression)
    assert t2["maxIter"][1] == 100 and t2["regParam"][1] == 0.0 and t2["family"][1] == "auto"
    assert all(len(v) == 3 for v in t2.values())
    sig = inspect.signature(b200fit.clustering.KMeans)
    assert all(p.default is not inspect.Parameter.empty for p in sig.parameters.values())


def test_evaluator_discovery_and_metric_names():
    ev = fw.list_evaluators(b200fit.evaluation)
    assert set(ev) == {"ClusteringEvaluator", "BinaryClassificationEvaluator", "MulticlassClassificationEvaluator"}
    for cls, metrics in ((b200fit.evaluation.ClusteringEvaluator, ["silhouette"]),
                         (b200fit.evaluation.BinaryClassificationEvaluator, ["areaUnderROC", "areaUnderPR"]),
                         (b200fit.evaluation.MulticlassClassificationEvaluator,
                          ["f1", "accuracy", "weightedPrecision", "weightedRecall"])):
        _, doc, table = fw.describe(cls)
        assert doc and all(len(v) == 3 and v[2] for v in table.values())
        assert table["metricName"][2].split("(")[-1].replace(")", "").split("|") == metrics
        assert hasattr(cls(), "apply") and cls().isLargerBetter()


def test_text_coercion_matches_gui_rules():
    assert fw.text_to_value("None") is None and fw.text_to_value("") is None
    assert fw.text_to_value("True") is True and fw.text_to_value("False") is False
    assert fw.text_to_value("8") == 8 and isinstance(fw.text_to_value("8"), int)
    assert fw.text_to_value("0.0") == 0.0 and isinstance(fw.text_to_value("0.0"), float)
    assert fw.text_to_value("1e-4") == 1e-4
    assert fw.text_to_value("k-means||") == "k-means||"


def test_param_map_is_resolved_by_name_and_types():
    km = b200fit.clustering.KMeans()
    pm = {fw.ForeignParam(km, "k", ""): 8, fw.ForeignParam(km, "tol", ""): 0, fw.ForeignParam(km, "seed", ""): None,
          fw.ForeignParam(km, "initMode", ""): "random"}
    c = km.copy(pm)
    assert c.getOrDefault("k") == 8 and c.getOrDefault("tol") == 0.0 and isinstance(c.getOrDefault("tol"), float)
    assert c.getOrDefault("seed") == km.getOrDefault("seed")  # None -> stays at default
    assert km.getOrDefault("k") == 2  # original untouched
    with pytest.raises(ValueError):
        km.copy({"nope": 1})
    with pytest.raises(TypeError):
        km.copy({"k": "many"})
    lr = b200fit.classification.LogisticRegression(regParam=1)
    assert lr.getRegParam() == 1.0 and lr.setMaxIter(5).getMaxIter() == 5


def test_dataframe_bridge_round_trip_with_a_pyspark_shaped_frame():
    """as_frame() collects a pyspark-like frame once through toPandas(); return_like() hands the result back as
    the caller's own frame type with only the new columns converted."""
    import pandas as pd

    from b200fit.frame import as_frame, return_like

    feats = [np.array([1.0, 2.0]), np.array([3.0, 4.0]), np.array([5.0, 6.0])]
    pdf = pd.DataFrame({"id": [7, 8, 9], "features": feats})
    sdf = fw.FakeSparkDataFrame(pdf)
    frame = as_frame(sdf)
    assert sdf.collected == 1 and frame.columns == ["id", "features"]
    assert frame.col("features").shape == (3, 2)  # vector cells stacked into one 2-D column
    out = frame.withColumn("prediction", np.array([0, 1, 1], dtype=np.int32))
    back = return_like(sdf, out)
    assert isinstance(back, fw.FakeSparkDataFrame) and back.columns == ["id", "features", "prediction"]
    res = back.toPandas()
    assert res["prediction"].tolist() == [0, 1, 1] and res["id"].tolist() == [7, 8, 9]
    assert all(np.array_equal(a, b) for a, b in zip(res["features"], feats))  # caller's column untouched
    # pandas in -> pandas out; vector-valued new columns become one array per row
    out2 = as_frame(pdf).withColumn("probability", np.array([[0.9, 0.1], [0.2, 0.8], [0.3, 0.7]]))
    back2 = return_like(pdf, out2)
    assert isinstance(back2, pd.DataFrame) and back2["probability"].iloc[1].tolist() == [0.2, 0.8]
    # our own frame type (and dicts) stay as they are
    cf = ColumnarFrame({"a": np.zeros(2)})
    assert return_like(cf, cf.withColumn("b", np.ones(2))).columns == ["a", "b"]


def test_thresholds_text_is_parsed_like_a_list():
    from b200fit.classification import _parse_thresholds

    assert _parse_thresholds(None) is None and _parse_thresholds("None") is None and _parse_thresholds("") is None
    assert _parse_thresholds("[0.5, 0.25, 0.25]") == [0.5, 0.25, 0.25]
    assert _parse_thresholds("0.5;0.5") == [0.5, 0.5] and _parse_thresholds((1, 2)) == [1.0, 2.0]
    for bad in ("[0.5, x]", [-0.1, 0.5], [0.0, 0.0, 1.0]):
        with pytest.raises(ValueError):
            _parse_thresholds(bad)


def test_constant_label_oracle_matches_mllib_short_circuit():
    rng = np.random.default_rng(0)
    X = rng.normal(size=(50, 3))
    r1 = oracle.lr_fit(X, np.ones(50, dtype=np.int64))
    r0 = oracle.lr_fit(X, np.zeros(50, dtype=np.int64))
    assert r1["intercept"] == np.inf and r0["intercept"] == -np.inf
    assert not r1["coefficients"].any() and r1["objectiveHistory"] == [] and r1["totalIterations"] == 0
    rm = oracle.lr_fit(X, np.full(50, 2, dtype=np.int64), family="multinomial")
    assert rm["intercept"].tolist() == [0.0, 0.0, np.inf] and rm["coefficients"].shape == (3, 3)
    # without an intercept there is no short circuit
    rn = oracle.lr_fit(X, np.ones(50, dtype=np.int64), fit_intercept=False, reg_param=0.1, max_iter=5)
    assert np.isfinite(rn["coefficients"]).all() and len(rn["objectiveHistory"]) > 1


def test_frame_behaves_like_the_dataframe_calls_the_widgets_make():
    df = ColumnarFrame({"b": np.arange(3), "a": np.array([1.5, 2.5, 3.5]), "cls": np.array([0, 1, 1])})
    assert sorted(df.columns) == ["a", "b", "cls"]
    out = df.withColumn("label", df["cls"].cast("double"))
    assert out.col("label").dtype == np.float64 and "label" not in df.columns
    assert out.count() == 3 and out.cache() is out
    with pytest.raises(ValueError):
        ColumnarFrame({"a": np.zeros(3), "b": np.zeros(4)})


def _toy_objective(seed, n=300, d=7, K=1, lam=0.03):
    rng = np.random.default_rng(seed)
    X = rng.normal(size=(n, d)) * rng.uniform(0.5, 2.0, size=d)
    if K == 1:
        y = (rng.uniform(size=n) < 1 / (1 + np.exp(-(X @ rng.normal(size=d))))).astype(float)
    else:
        y = np.argmax(X @ rng.normal(size=(d, K)) + rng.gumbel(size=(n, K)), axis=1)
    inv = np.ones(d)

    def f(x):
        if K == 1:
            loss, g, gb = oracle.lr_binary_lossgrad(X, y, None, x[:d], x[d], inv)
            return loss + 0.5 * lam * x[:d] @ x[:d], np.concatenate([g + lam * x[:d], [gb]])
        C = x[: K * d].reshape(K, d)
        loss, G, gb = oracle.lr_multinomial_lossgrad(X, y, None, C, x[K * d:], inv)
        return loss + 0.5 * lam * (C * C).sum(), np.concatenate([(G + lam * C).ravel(), gb])

    return f, np.zeros(K * d + K)


@pytest.mark.parametrize("seed,K", [(0, 1), (1, 1), (2, 3)])
def test_product_lbfgs_follows_the_oracle_trajectory(seed, K):
    f, x0 = _toy_objective(seed, K=K)
    mine = LBFGS(50, 10, 1e-9).run(f, x0)
    _, theirs = OLBFGS(50, 10, 1e-9).minimize(f, x0)
    assert len(mine) == len(theirs)
    for a, b in zip(mine, theirs):
        np.testing.assert_allclose(a.x, b["x"], rtol=0, atol=1e-12)
        assert a.adj_value == pytest.approx(b["adj_value"], abs=1e-13)
    assert mine[-1].n_evals == theirs[-1]["evals"]


@pytest.mark.parametrize("seed,K", [(3, 1), (4, 3)])
def test_product_owlqn_follows_the_oracle_trajectory(seed, K):
    f, x0 = _toy_objective(seed, K=K)
    d = (len(x0) - K) // K
    l1 = np.concatenate([np.full(K * d, 0.02), np.zeros(K)])
    mine = OWLQN(60, 10, l1, 1e-9).run(f, x0)
    _, theirs = OOWLQN(60, 10, l1, 1e-9).minimize(f, x0)
    assert len(mine) == len(theirs)
    for a, b in zip(mine, theirs):
        np.testing.assert_allclose(a.x, b["x"], rtol=0, atol=1e-12)
    assert (mine[-1].x[: K * d] == 0).any()
    assert mine[-1].adj_value < mine[0].adj_value


def test_product_has_no_path_into_the_oracle():
    pkg = os.path.join(ROOT, "orange3-spark_b200", "b200fit")
    for fn in os.listdir(pkg):
        if fn.endswith(".py"):
            src = open(os.path.join(pkg, fn)).read()
            assert not re.search(r"^\s*(from|import)\s+oracle\b", src, re.M), fn


def test_missing_gpu_fails_loudly():
    import torch

    if torch.cuda.is_available():
        pytest.skip("GPU present")
    with pytest.raises(_native.B200FitError):
        b200fit.clustering.KMeans(k=2, initialModel=np.zeros((2, 2))).fit(
            ColumnarFrame({"features": np.zeros((4, 2), dtype=np.float32)}))
