#!/usr/bin/env python
"""Golden vectors from REAL Apache Spark for the b200fit hot path -- run this OUTSIDE this repository's sandbox.

Why it exists.  The arithmetic behind the reference's Dataset Builder -> Estimator -> Model path lives in Spark
MLlib, which the reference neither vendors nor pins (setup.py:66-76) and which cannot run where this repository is
built (no JVM, no pyspark).  The repository's CPU oracle is therefore a restatement held to
three upstream doctest values only: "parity unpinned".  This script is the way out: anyone with `pip install
pyspark` and a JVM runs it once, commits the .npz files it writes, and tests/test_golden.py then holds BOTH the
oracle and the CUDA path to Spark's own per-iteration output (the slots are there and skip while the files are
absent).

    python tools/spark_oracle.py --out tests/golden [--cfg 1 2 3 4 5] [--master local[8]]

It needs pyspark and numpy and NOTHING from this repository: the data generator below is a copy of
oracle/synth.py (Philox4x32-10 keyed on (seed, global row, column), 8-term Irwin-Hall normal), the same generator
the CUDA library runs on the device (csrc/synth.cu), so all three sides see bit-identical fp32 inputs.

What is run (BASELINE.json configs at a reduced row count; d, k / K and every parameter as in the config):
  cfg 1  KMeans k=8      on 20 000 x 16,  mixture of 8,    maxIter 20, tol 1e-4   (widget defaults otherwise)
  cfg 2  KMeans k=1024   on 60 000 x 128, mixture of 1024, maxIter 5,  tol 0
  cfg 5  StandardScaler(withStd) -> KMeans k=256 on 40 000 x 64 with per-column scales, maxIter 6, tol 0
  cfg 3  LogisticRegression binary,      20 000 x 256, regParam 0.01, elasticNetParam 0,   tol 1e-6, maxIter 100
  cfg 4  LogisticRegression multinomial, 20 000 x 512, 20 classes, regParam 0.01, elasticNetParam 0.5
KMeans starts from FIXED centres (seeded distinct data rows) through pyspark.mllib's `initialModel` -- the only
upstream API that takes initial centres -- and is re-run with maxIterations = 1..T, which is deterministic, so the
file holds the centres after EVERY iteration.  LogisticRegression is re-run with maxIter = 1..T likewise
(coefficients after every iteration) and `summary.objectiveHistory` is taken from the longest run.

Every file records spark_version: the LR trajectory depends on it (Spark >= 3.2 centres the features when
fitIntercept is set; the optimum does not move, the path to it does).  The tests compare trajectories only for
versions whose aggregator the oracle restates (< 3.2) and converged values otherwise.
"""
import argparse
import os
import sys

import numpy as np

SEED = 20240611

# --------------------------------------------------------------------------------------------------------------
# Philox4x32-10 data generator (copy of oracle/synth.py; keep the two in step -- tests/test_golden.py checks it)
# --------------------------------------------------------------------------------------------------------------
M0, M1 = np.uint64(0xD2511F53), np.uint64(0xCD9E8D57)
W0, W1 = np.uint32(0x9E3779B9), np.uint32(0xBB67AE85)
IH_SCALE = np.float32(float.fromhex("0x1.3988e2p-16"))
_MASK = np.uint64(0xFFFFFFFF)


def philox4x32_10(c0, c1, c2, c3, k0, k1):
    c0, c1, c2, c3 = (np.asarray(v, dtype=np.uint32) for v in (c0, c1, c2, c3))
    c0, c1, c2, c3 = np.broadcast_arrays(c0, c1, c2, c3)
    k0, k1 = np.uint32(k0), np.uint32(k1)
    with np.errstate(over="ignore"):
        for _ in range(10):
            p0 = M0 * c0.astype(np.uint64)
            p1 = M1 * c2.astype(np.uint64)
            hi0, lo0 = (p0 >> np.uint64(32)).astype(np.uint32), (p0 & _MASK).astype(np.uint32)
            hi1, lo1 = (p1 >> np.uint64(32)).astype(np.uint32), (p1 & _MASK).astype(np.uint32)
            c0, c1, c2, c3 = hi1 ^ c1 ^ k0, lo1, hi0 ^ c3 ^ k1, lo0
            k0 = np.uint32((int(k0) + int(W0)) & 0xFFFFFFFF)
            k1 = np.uint32((int(k1) + int(W1)) & 0xFFFFFFFF)
    return c0, c1, c2, c3


def _ih_normal(words):
    s = np.zeros(words[0].shape, dtype=np.int64)
    for w in words:
        s += (w & np.uint32(0xFFFF)).astype(np.int64) + (w >> np.uint32(16)).astype(np.int64)
    return (s - 262140).astype(np.float32) * IH_SCALE


def znorm(seed, rows, cols, stream):
    rows = np.asarray(rows, dtype=np.uint64)
    r = philox4x32_10((rows & _MASK).astype(np.uint32), (rows >> np.uint64(32)).astype(np.uint32),
                      np.asarray(cols, dtype=np.uint32), np.uint32(stream), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)
    return _ih_normal(r)


def urand32(seed, rows, col, stream):
    rows = np.asarray(rows, dtype=np.uint64)
    return philox4x32_10((rows & _MASK).astype(np.uint32), (rows >> np.uint64(32)).astype(np.uint32),
                         np.uint32(col), np.uint32(stream), seed & 0xFFFFFFFF, (seed >> 32) & 0xFFFFFFFF)[0]


def mixture(n, row0, d, seed, n_components, center_scale, col_scale=None):
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    comp = (urand32(seed, rows, 0, 1) % np.uint32(n_components)).astype(np.uint64)
    cols = np.arange(d, dtype=np.uint32)
    z = znorm(seed, rows[:, None], cols[None, :], 0)
    mu = np.float32(center_scale) * znorm(seed, comp[:, None], cols[None, :], 2)
    X = (mu + z).astype(np.float32)
    if col_scale is not None:
        X = X * np.asarray(col_scale, dtype=np.float32)[None, :]
    return X


def logistic_features(n, row0, d, seed, col_scale=None):
    rows = np.arange(row0, row0 + n, dtype=np.uint64)
    z = znorm(seed, rows[:, None], np.arange(d, dtype=np.uint32)[None, :], 0)
    if col_scale is not None:
        z = z * np.asarray(col_scale, dtype=np.float32)[None, :]
    return z.astype(np.float32)


def logistic_labels(X, seed, w_true, b_true, n_classes, row0=0):
    """Labels from the uniform draw of stream 3, decided in fp64 on the fp32 features.  (The device generator
    decides them with device expf; these are stored in the golden file, so nothing has to reproduce them.)"""
    rows = np.arange(row0, row0 + len(X), dtype=np.uint64)
    u = (urand32(seed, rows, 0, 3).astype(np.float64) + 0.5) / 4294967296.0
    M = X.astype(np.float64) @ np.asarray(w_true, dtype=np.float64).T + np.asarray(b_true, dtype=np.float64)[None, :]
    if n_classes == 2:
        return (u < 1.0 / (1.0 + np.exp(-M[:, 0]))).astype(np.float64)
    P = np.exp(M - M.max(axis=1, keepdims=True))
    cdf = np.cumsum(P / P.sum(axis=1, keepdims=True), axis=1)
    return np.minimum((u[:, None] > cdf).sum(axis=1), n_classes - 1).astype(np.float64)


# --------------------------------------------------------------------------------------------------------------
# Spark runs
# --------------------------------------------------------------------------------------------------------------
def spark_session(master):
    from pyspark.sql import SparkSession

    return (SparkSession.builder.master(master).appName("b200fit-spark-oracle")
            .config("spark.ui.enabled", "false").config("spark.sql.shuffle.partitions", "8").getOrCreate())


def kmeans_golden(spark, X, C0, iters, tol):
    """Centres after each of `iters` Lloyd iterations from the fixed start C0, and the costs around them."""
    from pyspark.mllib.clustering import KMeans, KMeansModel
    from pyspark.mllib.linalg import Vectors

    rdd = spark.sparkContext.parallelize([Vectors.dense(r) for r in X.astype(np.float64)], 8).cache()
    init = KMeansModel([Vectors.dense(c) for c in C0])
    centres, cost_after = [], []
    for t in range(1, iters + 1):
        m = KMeans.train(rdd, len(C0), maxIterations=t, initialModel=init, epsilon=tol)
        centres.append(np.array([np.asarray(c) for c in m.clusterCenters]))
        cost_after.append(float(m.computeCost(rdd)))
        if t > 1 and np.array_equal(centres[-1], centres[-2]):
            break  # converged: further runs repeat the same model
    cost0 = float(init.computeCost(rdd))
    # cost_of_assignment[t] = cost measured with the centres that iteration t starts from (MLlib's trainingCost of a
    # run of t+1 iterations): [cost(C0), cost(C1), ...]
    return dict(centers_hist=np.stack(centres), cost_of_assignment=np.array([cost0] + cost_after[:-1]),
                cost_after_update=np.array(cost_after))


def lr_golden(spark, X, y, params, iters):
    from pyspark.ml.classification import LogisticRegression
    from pyspark.ml.linalg import Vectors

    df = spark.createDataFrame([(Vectors.dense(r), float(v)) for r, v in zip(X.astype(np.float64), y)],
                               ["features", "label"]).repartition(8).cache()
    coefs, icpts, hist, total = [], [], None, 0
    for t in range(1, iters + 1):
        m = LogisticRegression(maxIter=t, **params).fit(df)
        coefs.append(m.coefficientMatrix.toArray())
        icpts.append(np.asarray(m.interceptVector.toArray()))
        h = list(m.summary.objectiveHistory)
        if hist is not None and len(h) == len(hist):
            break  # the optimizer stopped before maxIter: converged
        hist, total = h, m.summary.totalIterations
    return dict(coef_hist=np.stack(coefs), icpt_hist=np.stack(icpts), objective_history=np.array(hist),
                total_iterations=total)


def run_cfg(spark, cfg, out_dir):
    import pyspark

    meta = dict(spark_version=pyspark.__version__, seed=SEED, cfg=cfg)
    if cfg in (1, 2, 5):
        n, d, k, iters, tol = {1: (20000, 16, 8, 20, 1e-4), 2: (60000, 128, 1024, 5, 0.0),
                               5: (40000, 64, 256, 6, 0.0)}[cfg]
        seed = SEED + (5 if cfg == 5 else 0)
        col = None
        if cfg == 5:
            col = np.exp(np.random.default_rng(seed).uniform(np.log(0.1), np.log(10.0), size=d)).astype(np.float32)
        X = mixture(n, 0, d, seed, k, 2.0 if cfg != 1 else 3.0, col)
        rows0 = np.random.default_rng(2).choice(n, size=k, replace=False)
        extra = {}
        if cfg == 5:
            from pyspark.ml.feature import StandardScaler
            from pyspark.ml.linalg import Vectors

            df = spark.createDataFrame([(Vectors.dense(r),) for r in X.astype(np.float64)], ["raw"])
            sm = StandardScaler(inputCol="raw", outputCol="features", withMean=False, withStd=True).fit(df)
            extra["scaler_std"] = np.asarray(sm.std.toArray())
            Xs = np.array([np.asarray(r.features.toArray()) for r in sm.transform(df).select("features").collect()])
            extra["col_scale"] = col
        else:
            Xs = X.astype(np.float64)
        C0 = Xs[rows0]
        res = kmeans_golden(spark, Xs, C0, iters, tol)
        np.savez_compressed(os.path.join(out_dir, "spark_cfg%d.npz" % cfg), n=n, d=d, k=k, tol=tol, data_seed=seed,
                            center_scale=2.0 if cfg != 1 else 3.0, init_rows=rows0, **extra, **res,
                            **{k_: np.array(v) for k_, v in meta.items()})
    else:
        n, d, K = {3: (20000, 256, 2), 4: (20000, 512, 20)}[cfg]
        seed = SEED + (1 if cfg == 3 else 2)
        rng = np.random.default_rng(seed)
        scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
        w_true = (rng.standard_normal((1 if K == 2 else K, d)) / np.sqrt(d)).astype(np.float32)
        b_true = np.array([0.3], dtype=np.float32) if K == 2 else np.zeros(K, dtype=np.float32)
        X = logistic_features(n, 0, d, seed, scale)
        y = logistic_labels(X, seed, w_true, b_true, K)
        params = dict(regParam=0.01, elasticNetParam=0.0 if cfg == 3 else 0.5, tol=1e-6, fitIntercept=True,
                      standardization=True, family="binomial" if cfg == 3 else "multinomial")
        res = lr_golden(spark, X, y, params, 100 if cfg == 3 else 60)
        np.savez_compressed(os.path.join(out_dir, "spark_cfg%d.npz" % cfg), n=n, d=d, K=K, data_seed=seed,
                            col_scale=scale, y=y, reg_param=0.01, elastic_net=params["elasticNetParam"], **res,
                            **{k_: np.array(v) for k_, v in meta.items()})
    print("cfg %d written" % cfg)


def main():
    ap = argparse.ArgumentParser(description=__doc__.split("\n")[0])
    ap.add_argument("--out", default="tests/golden")
    ap.add_argument("--cfg", type=int, nargs="*", default=[1, 2, 3, 4, 5])
    ap.add_argument("--master", default="local[*]")
    a = ap.parse_args()
    try:
        import pyspark  # noqa: F401
    except ImportError:
        sys.exit("pyspark is not importable here; run this script where Spark is installed (pip install pyspark + a JVM)")
    os.makedirs(a.out, exist_ok=True)
    spark = spark_session(a.master)
    try:
        for cfg in a.cfg:
            run_cfg(spark, cfg, a.out)
    finally:
        spark.stop()


if __name__ == "__main__":
    main()
