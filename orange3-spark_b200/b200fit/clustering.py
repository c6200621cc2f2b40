"""b200fit.clustering -- KMeans / KMeansModel with the pyspark.ml.clustering surface.

Swap point in the reference: `module = clustering`
(orangecontrib/spark/widgets/ml/spark_ml_clustering.py:14); the widget then calls
`KMeans().fit(in_df, params=paramMap)` (orangecontrib/spark/base/spark_ml_estimator.py:19-25)
and later `model.transform(in_df)` (orangecontrib/spark/widgets/ml/spark_ml_model.py:53).
Semantics follow MLlib's Lloyd iteration; the arithmetic runs in
libb200fit.so (b2f_kmeans_*), rows sharded over ranks with one all-reduce per iteration.
"""
import numpy as np

from . import _native as N
from .base import Estimator, Model
from .dist import Comm
from .feature import to_device_matrix, to_device_vector
from .frame import as_frame
from .param import TypeConverters as TC
from .param import make_accessors


def _torch():
    import torch

    return torch


class CudaLloydEngine:
    """Device state + kernel calls for one rank's row shard."""

    def __init__(self, X, k, w=None, algo=N.KM_AUTO, ws_bytes=None):
        N.require_cuda()
        torch = _torch()
        self.L = N.lib()
        self.X, self.k, self.d, self.w, self.algo = X, int(k), X.d, w, algo
        dev = X.data.device
        self.dev = dev
        self.centers = torch.zeros((self.k, self.d), dtype=torch.float64, device=dev)
        self.prep = torch.empty(self.L.b2f_kmeans_prep_bytes(self.k, self.d), dtype=torch.uint8, device=dev)
        self.acc = torch.empty(self.L.b2f_kmeans_acc_bytes(self.k, self.d), dtype=torch.uint8, device=dev)
        self.plen = int(self.L.b2f_kmeans_payload_len(self.k, self.d))
        self.payload = self.acc[: self.plen * 8].view(torch.float64)
        self.status = torch.zeros(2, dtype=torch.int32, device=dev)
        self.status_f64 = torch.zeros(2, dtype=torch.float64, device=dev)
        self.launches = 0
        self.fallbacks = 0
        self._xmax = None
        # low-precision filter pass + candidate check (at most four centres, fp32 / fp64) + 3xTF32 re-check (B2F_KM_TCH fp16 / B2F_KM_TCB bf16x3; AUTO = TCH)
        # needs a labels buffer and a workspace.  Decided from the SHAPE only (never from this rank's row count):
        # every rank must take the same branches between collectives, also one whose shard is empty.
        self.use_ws = (algo in (N.KM_AUTO, N.KM_TCB, N.KM_TCH) and w is None and self.d in (64, 128)
                       and self.k % 128 == 0 and self.k <= 65536 and X.ld % 4 == 0)
        if algo in (N.KM_TCB, N.KM_TCH) and not self.use_ws:
            raise ValueError("B2F_KM_TCB / B2F_KM_TCH need d in {64, 128}, k % 128 == 0 and no row weights")
        self.col_scale = getattr(X, "col_scale", None)
        if self.col_scale is not None and not (self.use_ws and algo in (N.KM_AUTO, N.KM_TCH)):
            raise ValueError("a scaled view needs the fp16 filter pipeline; materialise it first")
        if self.use_ws:
            # (ws_bytes: tests shrink the workspace to force the overflow fallback)
            self.ws_bytes = int(self.L.b2f_kmeans_ws_bytes(X.n, self.k, self.d)) if ws_bytes is None else int(ws_bytes)
            self.ws = torch.zeros(self.ws_bytes, dtype=torch.uint8, device=dev)
            self.labels_buf = torch.empty(X.n, dtype=torch.int32, device=dev)

    def set_centers(self, C):
        torch = _torch()
        C = np.ascontiguousarray(np.asarray(C, dtype=np.float64))
        if C.shape != (self.k, self.d):
            raise ValueError("initial centres have shape %s, expected %s" % (C.shape, (self.k, self.d)))
        self.centers.copy_(torch.from_numpy(C))
        self._prepare()

    def _x_norm_max(self):
        """Largest row norm of this shard (of the scaled rows for a view), once: lets the fp16 filter use one key
        scale for all rows (b2f_kmeans_prepare_x).  Every rank may use its own value: the scale only has to cover the
        rows this rank's kernels meet."""
        if self._xmax is None:
            self._xmax = 0.0
            if self.use_ws and self.X.n > 0:
                out = _torch().zeros(1, dtype=_torch().float32, device=self.dev)
                N.check(self.L.b2f_row_norm_max(N.ptr(self.X.data), self.X.n, self.d, self.X.ld,
                                                N.ptr(self.col_scale), N.ptr(out), N.stream_ptr()), "b2f_row_norm_max")
                self._xmax = float(out.item())
                self.launches += 1
        return self._xmax

    def _prepare(self):
        N.check(self.L.b2f_kmeans_prepare_x(N.ptr(self.centers), self.k, self.d, self._x_norm_max(), N.ptr(self.prep),
                                            N.stream_ptr()), "b2f_kmeans_prepare_x")
        self.launches += 1

    def step(self, labels=None, exact=False, events=None):
        """Zero accumulators, assign + accumulate this shard, fold -> fp64 payload (device).
        exact=True (used when a list of the filter pass overflowed): 3xTF32 over every row WITH its certainty
        test and the fp64 tail -- the same label guarantee as the filter pipeline, just slower."""
        s = N.stream_ptr()
        X = self.X
        N.check(self.L.b2f_kmeans_acc_zero(N.ptr(self.acc), self.k, self.d, s), "b2f_kmeans_acc_zero")
        if events:
            events[0].record()
        if self.use_ws:
            lab = labels if labels is not None else self.labels_buf
            if exact and self.col_scale is not None:
                # rare path (list overflow on a scaled view): run it on the materialised copy
                Xm = X.materialized()
                N.check(self.L.b2f_kmeans_step_ws(N.ptr(Xm.data), Xm.n, Xm.d, Xm.ld, None, N.ptr(self.prep), self.k,
                                                  N.ptr(self.acc), N.ptr(lab), N.ptr(self.ws), self.ws_bytes,
                                                  N.KM_TC3, s), "b2f_kmeans_step_ws")
                self.launches += 3
            elif self.col_scale is not None:
                N.check(self.L.b2f_kmeans_step_scaled(N.ptr(X.data), X.n, X.d, X.ld, N.ptr(self.col_scale),
                                                      N.ptr(self.prep), self.k, N.ptr(self.acc), N.ptr(lab),
                                                      N.ptr(self.ws), self.ws_bytes, s), "b2f_kmeans_step_scaled")
                self.launches += 10  # as below + the two folds around the filter kernel
            else:
                algo = N.KM_TC3 if exact else self.algo
                N.check(self.L.b2f_kmeans_step_ws(N.ptr(X.data), X.n, X.d, X.ld, None, N.ptr(self.prep), self.k,
                                                  N.ptr(self.acc), N.ptr(lab), N.ptr(self.ws), self.ws_bytes, algo,
                                                  s), "b2f_kmeans_step_ws")
                # TCH: 2 cost-term kernels + filter + candidate check + gather + 3xTF32 re-check + scatter + exact tail;
                # TCB: filter + gather + re-check + scatter + exact tail; exact: set-cap + 3xTF32 + exact tail
                self.launches += 3 if exact else (5 if self.algo == N.KM_TCB else 8)
        else:
            N.check(self.L.b2f_kmeans_step(N.ptr(X.data), X.n, X.d, X.ld, N.ptr(self.w), N.ptr(self.prep), self.k,
                                           N.ptr(self.acc), N.ptr(labels), self.algo, s), "b2f_kmeans_step")
            self.launches += 1
        if events:
            events[1].record()
        N.check(self.L.b2f_kmeans_fold(N.ptr(self.acc), self.k, self.d, s), "b2f_kmeans_fold")
        self.launches += 1
        return self.payload

    def rows_counted(self, payload):
        """Rows that were accumulated (last payload entry); below the row total means the re-check list
        of the filter pass overflowed and the pass must be redone with exact=True."""
        return float(payload[self.k * self.d + self.k + 1].item())

    def finalize_async(self, payload, tol):
        s = N.stream_ptr()
        N.check(self.L.b2f_kmeans_finalize(N.ptr(payload), self.k, self.d, float(tol), N.ptr(self.centers),
                                           N.ptr(self.status), N.ptr(self.status_f64), s), "b2f_kmeans_finalize")
        self.launches += 2
        self._prepare()

    def finalize(self, payload, tol):
        """Centre update from the reduced payload; returns (converged, cost of this assignment)."""
        self.finalize_async(payload, tol)
        st = self.status.cpu()
        sf = self.status_f64.cpu()
        return bool(st[0].item() == 1), float(sf[1].item())

    def get_centers(self):
        return self.centers.cpu().numpy().copy()

    def get_counts(self, payload):
        kd = self.k * self.d
        return payload[kd: kd + self.k].cpu().numpy().copy()


def lloyd(engine, comm, max_iter, tol, n_total=None):
    """A.2 iteration loop: stop at maxIter or when every moved centre moved <= tol; the final
    model holds the updated centres; trainingCost is the cost of the LAST assignment step.
    Every branch that precedes a collective is taken on all-reduced values, so all ranks stay in step."""
    it, converged, costs = 0, False, []
    payload = None
    use_ws = bool(getattr(engine, "use_ws", False))  # shape-derived: identical on every rank
    if use_ws and n_total is None:
        import torch

        t = torch.tensor([float(engine.X.n)], dtype=torch.float64, device=engine.dev)
        comm.all_reduce_sum(t)
        n_total = float(t.item())
    while it < max_iter and not converged:
        payload = engine.step()
        comm.all_reduce_sum(payload)
        if use_ws and engine.rows_counted(payload) != n_total:
            # some rank dropped uncertain rows (a list of the filter pass overflowed): redo the pass on every rank
            # with 3xTF32 + certainty test + fp64 tail, whose list has room for every row
            engine.fallbacks += 1
            payload = engine.step(exact=True)
            comm.all_reduce_sum(payload)
            if engine.rows_counted(payload) != n_total:
                raise RuntimeError("KMeans: %d of %d rows were not assigned (workspace smaller than "
                                   "b2f_kmeans_ws_bytes?)" % (n_total - engine.rows_counted(payload), n_total))
        converged, cost = engine.finalize(payload, tol)
        costs.append(cost)
        it += 1
    return it, costs, converged, payload


def ws_dropped(ws):
    """Rows a b2f_kmeans_*_ws call could not place in its lists (header of the workspace, include/b200fit.h)."""
    h = ws[:16].view(_torch().int32).cpu().numpy()
    return max(0, int(h[0]) - int(h[3])) + max(0, int(h[2]) - int(h[3]))


def predict_labels(X, centers, want_dist=False, algo=N.KM_AUTO, info=None):
    """Assignment only (KMeansModel.transform / computeCost / initialisers).  Where the shape allows it the labels
    come from the same certainty pipeline as the fit (filter pass, candidate check, 3xTF32 re-check, fp64 tail), so they
    equal the fp64 assignment except for fp64-level ties; distances are those to the labelled centre."""
    N.require_cuda()
    torch = _torch()
    L = N.lib()
    C = np.ascontiguousarray(np.asarray(centers, dtype=np.float64))
    k, d = C.shape
    if d != X.d:
        raise ValueError("model has %d features, data has %d" % (d, X.d))
    dev = X.data.device
    cdev = torch.from_numpy(C).to(dev)
    prep = torch.empty(L.b2f_kmeans_prep_bytes(k, d), dtype=torch.uint8, device=dev)
    N.check(L.b2f_kmeans_prepare(N.ptr(cdev), k, d, N.ptr(prep), N.stream_ptr()), "b2f_kmeans_prepare")
    labels = torch.empty(X.n, dtype=torch.int32, device=dev)
    dist = torch.empty(X.n, dtype=torch.float32, device=dev) if want_dist else None
    tc_shape = d % 32 == 0 and 32 <= d <= 128 and k % 128 == 0 and k <= 65536 and X.ld % 4 == 0
    scale = getattr(X, "col_scale", None)
    if scale is not None and not (tc_shape and d in (64, 128) and algo in (N.KM_AUTO, N.KM_TCH)):
        X, scale = X.materialized(), None
    if not tc_shape or algo in (N.KM_GENERIC, N.KM_TC1) or X.n == 0:
        N.check(L.b2f_kmeans_predict(N.ptr(X.data), X.n, d, X.ld, N.ptr(prep), k, N.ptr(labels), N.ptr(dist), algo,
                                     N.stream_ptr()), "b2f_kmeans_predict")
        return (labels, dist) if want_dist else labels
    ws_bytes = int(L.b2f_kmeans_ws_bytes(X.n, k, d))
    ws = torch.zeros(ws_bytes, dtype=torch.uint8, device=dev)
    N.check(L.b2f_kmeans_predict_ws(N.ptr(X.data), X.n, d, X.ld, N.ptr(scale), N.ptr(prep), k, N.ptr(labels),
                                    N.ptr(dist), N.ptr(ws), ws_bytes, algo, N.stream_ptr()), "b2f_kmeans_predict_ws")
    dropped = ws_dropped(ws)
    if info is not None:
        info["dropped_first_pass"] = dropped
    if dropped:  # a list overflowed: 3xTF32 over every row with its certainty test + fp64 tail (room for all rows)
        Xm = X.materialized() if scale is not None else X
        N.check(L.b2f_kmeans_predict_ws(N.ptr(Xm.data), Xm.n, d, Xm.ld, None, N.ptr(prep), k, N.ptr(labels),
                                        N.ptr(dist), N.ptr(ws), ws_bytes, N.KM_TC3, N.stream_ptr()),
                "b2f_kmeans_predict_ws")
        if ws_dropped(ws):
            raise RuntimeError("KMeans predict: rows left unassigned after the exact fallback")
    return (labels, dist) if want_dist else labels


_DEFAULT_SEED = int.from_bytes(b"KMea", "big") & 0x7FFFFFFF  # fixed: ranks must agree

_KM_PARAMS = (
    ("featuresCol", "features column name.", TC.toString),
    ("predictionCol", "prediction column name.", TC.toString),
    ("k", "The number of clusters to create. Must be > 1.", TC.toInt),
    ("initMode", "The initialization algorithm. This can be either \"random\" to choose random points as "
                 "initial cluster centers, or \"k-means||\" to use a parallel variant of k-means++", TC.toString),
    ("initSteps", "The number of steps for k-means|| initialization mode. Must be > 0.", TC.toInt),
    ("tol", "the convergence tolerance for iterative algorithms (>= 0).", TC.toFloat),
    ("maxIter", "max number of iterations (>= 0).", TC.toInt),
    ("seed", "random seed.", TC.toInt),
    ("distanceMeasure", "the distance measure. Supported options: 'euclidean'.", TC.toString),
    ("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
     TC.toString),
    ("solver", "The solver algorithm for optimization. Supported options: 'auto', 'row', 'block' "
               "(all map to the fused B200 kernel).", TC.toString),
    ("maxBlockSizeInMB", "accepted for pyspark compatibility; ignored.", TC.toFloat),
    ("initialModel", "extension: a KMeansModel or a [k, d] array of initial cluster centers; when set, "
                     "initMode / initSteps / seed are ignored.", TC.identity),
)


class KMeansSummary:
    """trainingCost = cost of the last assignment step; numIter = iterations executed (A.2)."""

    def __init__(self, k, num_iter, training_cost, cluster_sizes, cost_history, converged):
        self.k = k
        self.numIter = num_iter
        self.trainingCost = training_cost
        self.clusterSizes = [int(round(c)) for c in cluster_sizes]
        self.costHistory = list(cost_history)
        self.converged = converged


@make_accessors
class KMeansModel(Model):
    """Model fitted by KMeans."""

    _declared = tuple(p for p in _KM_PARAMS if p[0] in ("featuresCol", "predictionCol", "k", "distanceMeasure"))

    def __init__(self, centers=None, summary=None):
        super().__init__()
        self._setDefault(featuresCol="features", predictionCol="prediction", k=2, distanceMeasure="euclidean")
        self._centers = None if centers is None else np.ascontiguousarray(np.asarray(centers, dtype=np.float64))
        self._summary = summary

    def clusterCenters(self):
        return [c.copy() for c in self._centers]

    @property
    def hasSummary(self):
        return self._summary is not None

    @property
    def summary(self):
        if self._summary is None:
            raise RuntimeError("No training summary available for this KMeansModel")
        return self._summary

    def _transform(self, dataset):
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("featuresCol")), allow_scaled=True)
        labels = predict_labels(X, self._centers)
        return frame.withColumn(self.getOrDefault("predictionCol"), labels)

    def predict(self, value):
        v = np.asarray(value, dtype=np.float64).reshape(1, -1)
        return int(((self._centers - v) ** 2).sum(1).argmin())

    def computeCost(self, dataset):
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("featuresCol")), allow_scaled=True)
        _, dist = predict_labels(X, self._centers, want_dist=True)
        c = dist.double().sum().reshape(1)
        Comm().all_reduce_sum(c)
        return float(c.item())


@make_accessors
class KMeans(Estimator):
    """K-means clustering with a k-means++ like initialization mode (the k-means|| algorithm
    by Bahmani et al), Lloyd iterations fused into one B200 kernel per pass.
    """

    _declared = _KM_PARAMS

    def __init__(self, featuresCol="features", predictionCol="prediction", k=2, initMode="k-means||", initSteps=2,
                 tol=1e-4, maxIter=20, seed=None, distanceMeasure="euclidean", weightCol=None, solver="auto",
                 maxBlockSizeInMB=0.0, initialModel=None):
        super().__init__()
        self._setDefault(featuresCol="features", predictionCol="prediction", k=2, initMode="k-means||", initSteps=2,
                         tol=1e-4, maxIter=20, seed=_DEFAULT_SEED,
                         distanceMeasure="euclidean", solver="auto", maxBlockSizeInMB=0.0)
        self._set(featuresCol=featuresCol, predictionCol=predictionCol, k=k, initMode=initMode, initSteps=initSteps,
                  tol=tol, maxIter=maxIter, seed=seed, distanceMeasure=distanceMeasure, weightCol=weightCol,
                  solver=solver, maxBlockSizeInMB=maxBlockSizeInMB, initialModel=initialModel)
        self._algo = N.KM_AUTO

    def setParams(self, **kw):
        return self._set(**kw)

    def _fit(self, dataset):
        from ._kmeans_init import init_kmeans_parallel, init_random

        torch = _torch()
        g = self.getOrDefault
        k, max_iter, tol = g("k"), g("maxIter"), g("tol")
        if k is None or k < 1:
            raise ValueError("KMeans: k must be >= 1, got %r" % (k,))
        if max_iter < 0 or tol < 0:
            raise ValueError("KMeans: maxIter and tol must be >= 0")
        if g("distanceMeasure") != "euclidean":
            raise ValueError("KMeans: only distanceMeasure='euclidean' is supported, got %r" % g("distanceMeasure"))
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(g("featuresCol")), allow_scaled=True)
        wcol = g("weightCol")
        w = to_device_vector(frame.col(wcol), X.n) if wcol else None
        if getattr(X, "col_scale", None) is not None and not (
                w is None and self._algo in (N.KM_AUTO, N.KM_TCH)
                and N.lib().b2f_kmeans_scaled_supported(int(k), X.d, X.ld)):
            X = X.materialized()  # shapes the fused scaled pass does not take: write the scaled copy once
        comm = Comm()
        nt = torch.tensor([X.n], dtype=torch.int64, device=X.data.device)
        comm.all_reduce_sum(nt)
        n_total = int(nt.item())
        init = g("initialModel")
        if init is not None:
            C0 = np.asarray(init.clusterCenters() if hasattr(init, "clusterCenters") else init, dtype=np.float64)
            if C0.ndim != 2 or C0.shape[0] != k or C0.shape[1] != X.d:
                raise ValueError("KMeans: initialModel has shape %s, expected (%d, %d)" % (C0.shape, k, X.d))
        elif g("initMode") == "random":
            C0 = init_random(X, n_total, k, g("seed"), comm)
        elif g("initMode") == "k-means||":
            if g("initSteps") <= 0:
                raise ValueError("KMeans: initSteps must be > 0")
            C0 = init_kmeans_parallel(X, n_total, k, g("seed"), g("initSteps"), comm,
                                      lambda X_, C_, wd: predict_labels(X_, C_, want_dist=wd))
        else:
            raise ValueError("KMeans: unsupported initMode %r" % g("initMode"))
        eng = CudaLloydEngine(X, k, w=w, algo=self._algo)
        eng.set_centers(C0)
        it, costs, converged, payload = lloyd(eng, comm, max_iter, tol, float(n_total))
        sizes = eng.get_counts(payload) if payload is not None else np.zeros(k)
        model = KMeansModel(eng.get_centers(),
                            KMeansSummary(k, it, costs[-1] if costs else float("nan"), sizes, costs, converged))
        model._set(featuresCol=g("featuresCol"), predictionCol=g("predictionCol"), k=k,
                   distanceMeasure=g("distanceMeasure"))
        return model
