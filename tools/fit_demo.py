"""Full-size fits through the public API (one GPU): BASELINE configs 2, 3 and 4.

  python tools/fit_demo.py [--scale 1.0]

Data are generated on the device (b2f_synth_*), wrapped in a ColumnarFrame and handed to
KMeans / LogisticRegression exactly as the widgets would; prints one JSON line per config."""
import argparse
import json
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, os.path.join(ROOT, "orange3-spark_b200"))
import numpy as np
import torch

from b200fit import _native as N
from b200fit._kmeans_init import init_random
from b200fit.classification import LogisticRegression
from b200fit.clustering import KMeans
from b200fit.dist import Comm
from b200fit.frame import ColumnarFrame, DeviceMatrix

ap = argparse.ArgumentParser()
ap.add_argument("--scale", type=float, default=1.0, help="fraction of the BASELINE row counts")
ap.add_argument("--only", default="", help="comma list out of kmeans,lr,lrm")
a = ap.parse_args()
only = set(a.only.split(",")) if a.only else {"kmeans", "lr", "lrm"}
L = N.lib()
dev = torch.device("cuda", 0)
SEED = 20240611


def timed(fn):
    torch.cuda.synchronize()
    t0 = time.perf_counter()
    out = fn()
    torch.cuda.synchronize()
    return out, time.perf_counter() - t0


if "kmeans" in only:
    n, d, k = int(100_000_000 * a.scale), 128, 1024
    X = torch.empty((n, d), dtype=torch.float32, device=dev)
    N.check(L.b2f_synth_mixture(N.ptr(X), n, 0, d, d, SEED, k, 2.0, None, N.stream_ptr()), "synth")
    dm = DeviceMatrix(X, d)
    C0 = init_random(dm, n, k, SEED, Comm())
    model, dt = timed(lambda: KMeans(k=k, maxIter=20, tol=0.0, initialModel=C0).fit(ColumnarFrame({"features": dm})))
    s = model.summary
    print(json.dumps({"config": "KMeans k=1024 maxIter=20 on %dx128" % n, "seconds": dt, "iterations": s.numIter,
                      "iter_per_s": s.numIter / dt, "cost_first": s.costHistory[0], "cost_last": s.trainingCost,
                      "cost_monotone": all(s.costHistory[i + 1] <= s.costHistory[i] * (1 + 1e-9)
                                           for i in range(len(s.costHistory) - 1)),
                      "sum_sizes": sum(s.clusterSizes), "empty_clusters": sum(1 for c in s.clusterSizes if c == 0)}))
    del X, dm, model
    torch.cuda.empty_cache()


def lr_data(n, d, K, seed):
    rng = np.random.default_rng(seed)
    scale = np.exp(rng.uniform(np.log(0.5), np.log(2.0), size=d)).astype(np.float32)
    KM = 1 if K == 2 else K
    w_true = (rng.standard_normal((KM, d)) / np.sqrt(d)).astype(np.float32)
    b_true = np.full(KM, 0.3 if K == 2 else 0.0, dtype=np.float32)
    X = torch.empty((n, d), dtype=torch.float32, device=dev)
    y = torch.empty(n, dtype=torch.float32, device=dev)
    sc_d, w_d, b_d = (torch.from_numpy(t).to(dev) for t in (scale, w_true, b_true))  # keep alive across the call
    N.check(L.b2f_synth_logistic(N.ptr(X), N.ptr(y), n, 0, d, d, seed, N.ptr(sc_d), N.ptr(w_d), N.ptr(b_d), K,
                                 N.stream_ptr()), "synth")
    torch.cuda.synchronize()
    return X, y, w_true


if "lr" in only:
    n, d = int(50_000_000 * a.scale), 256
    X, y, w_true = lr_data(n, d, 2, SEED + 1)
    frame = ColumnarFrame({"features": DeviceMatrix(X, d), "label": y})
    model, dt = timed(lambda: LogisticRegression(regParam=0.01, maxIter=100, tol=1e-6).fit(frame))
    s = model.summary
    cos = float(np.dot(model.coefficients, w_true[0]) / (np.linalg.norm(model.coefficients) * np.linalg.norm(w_true[0])))
    print(json.dumps({"config": "binary LogisticRegression L-BFGS regParam=0.01 on %dx256" % n, "seconds": dt,
                      "iterations": s.totalIterations, "lossgrad_passes": s.lossGradPasses,
                      "passes_per_s_incl_host": s.lossGradPasses / dt, "objective_first": s.objectiveHistory[0],
                      "objective_last": s.objectiveHistory[-1], "objective_history": s.objectiveHistory,
                      "passes_at_state": s.passesAtState, "cosine_to_true_weights": cos,
                      "intercept": model.intercept}))
    del X, y, frame, model
    torch.cuda.empty_cache()

if "lrm" in only:
    n, d, K = int(20_000_000 * a.scale), 512, 20
    X, y, w_true = lr_data(n, d, K, SEED + 2)
    frame = ColumnarFrame({"features": DeviceMatrix(X, d), "label": y})
    # with regParam = 0.01 the L1 part (0.005) exceeds every |gradient entry| at the start for this data and
    # OWL-QN correctly stops at the all-zero model after one pass; 0.001 exercises the optimizer
    model, dt = timed(lambda: LogisticRegression(regParam=0.001, elasticNetParam=0.5, family="multinomial",
                                                 maxIter=100, tol=1e-6).fit(frame))
    s = model.summary
    print(json.dumps({"config": "multinomial LogisticRegression (20 classes, regParam 0.001, elasticNet 0.5, OWL-QN) "
                                "on %dx512" % n,
                      "seconds": dt, "iterations": s.totalIterations, "lossgrad_passes": s.lossGradPasses,
                      "passes_per_s_incl_host": s.lossGradPasses / dt, "objective_first": s.objectiveHistory[0],
                      "objective_last": s.objectiveHistory[-1],
                      "zero_coefficients": int((model.coefficientMatrix == 0).sum()),
                      "n_coefficients": int(model.coefficientMatrix.size)}))
