"""b200fit.classification -- LogisticRegression with the pyspark.ml.classification surface.

Swap point in the reference: `module = classification`
(orangecontrib/spark/widgets/ml/spark_ml_classification.py:15); the widget calls
`LogisticRegression().fit(in_df, params=paramMap)`
(orangecontrib/spark/base/spark_ml_estimator.py:19-25) and `model.transform(in_df)`
(orangecontrib/spark/widgets/ml/spark_ml_model.py:53).  Objective, standardisation,
initial point and post-processing follow MLlib; every loss+gradient
evaluation is one fused GPU pass (b2f_lr_*_lossgrad) and one all-reduce.
"""
import numpy as np

from . import _native as N
from .base import Estimator, Model
from .dist import Comm
from .feature import column_stats, to_device_matrix, to_device_vector
from .frame import as_frame
from .optim import LBFGS, OWLQN
from .param import TypeConverters as TC
from .param import make_accessors


def _torch():
    import torch

    return torch


class CudaLogRegEngine:
    """One rank's shard: X, y, w resident in HBM; lossgrad() returns the fp64 payload
    [K*d raw gradient | K intercept gradient | loss sum | weight sum] (device)."""

    def __init__(self, X, y, w, n_classes):
        N.require_cuda()
        torch = _torch()
        self.L = N.lib()
        self.X, self.y, self.w = X, y, w
        self.nc = int(n_classes)
        self.K = 1 if self.nc == 2 else self.nc
        self.d = X.d
        dev = X.data.device
        self.plen = self.K * self.d + self.K + 2
        self.out = torch.zeros(self.plen, dtype=torch.float64, device=dev)
        self.scratch = torch.empty(max(1, self.L.b2f_lr_scratch_bytes(self.d, self.nc)), dtype=torch.uint8, device=dev)
        self.beta = torch.zeros(self.K * self.d, dtype=torch.float64, device=dev)
        self.icpt = torch.zeros(self.K, dtype=torch.float64, device=dev)
        # pinned staging for the per-pass traffic of a fit: iterate up (K*d + K doubles), payload down
        self.h_x = torch.empty(self.K * self.d + self.K, dtype=torch.float64, pin_memory=True)
        self.h_out = torch.empty(self.plen, dtype=torch.float64, pin_memory=True)
        self.launches = 0

    def lossgrad_device(self):
        """Uses the beta / icpt already on the device; leaves the payload in self.out."""
        X, s = self.X, N.stream_ptr()
        self.out.zero_()
        if self.K == 1:
            N.check(self.L.b2f_lr_binary_lossgrad(N.ptr(X.data), N.ptr(self.y), N.ptr(self.w), X.n, X.d, X.ld,
                                                  N.ptr(self.beta), N.ptr(self.icpt), N.ptr(self.out),
                                                  N.ptr(self.scratch), s), "b2f_lr_binary_lossgrad")
        else:
            N.check(self.L.b2f_lr_multinomial_lossgrad(N.ptr(X.data), N.ptr(self.y), N.ptr(self.w), X.n, X.d, X.ld,
                                                       self.nc, N.ptr(self.beta), N.ptr(self.icpt), N.ptr(self.out),
                                                       N.ptr(self.scratch), s), "b2f_lr_multinomial_lossgrad")
        self.launches += 2
        return self.out

    def lossgrad(self, beta, icpt):
        kd = self.K * self.d
        hx = self.h_x.numpy()
        hx[:kd] = np.asarray(beta, dtype=np.float64).ravel()
        hx[kd:] = np.asarray(icpt, dtype=np.float64).ravel()
        self.beta.copy_(self.h_x[:kd], non_blocking=True)
        self.icpt.copy_(self.h_x[kd:], non_blocking=True)
        return self.lossgrad_device()

    def payload_to_host(self, payload):
        """The (all-reduced) payload as a numpy view of pinned memory: one async copy + one stream sync."""
        self.h_out.copy_(payload, non_blocking=True)
        _torch().cuda.current_stream().synchronize()
        return self.h_out.numpy()


def make_objective(engine, comm, K, d, inv_std, l2w, fit_intercept, active):
    """Objective in standardized space: x = [coef (K*d, row-major) | intercepts (K)]."""

    def f(x):
        C = x[: K * d].reshape(K, d) * active[None, :]
        b = x[K * d:] if fit_intercept else np.zeros(K)
        payload = engine.lossgrad(C * inv_std[None, :], b)
        comm.all_reduce_sum(payload)
        p = engine.payload_to_host(payload) if hasattr(engine, "payload_to_host") else payload.cpu().numpy()
        W = p[K * d + K + 1]
        G = p[: K * d].reshape(K, d) * inv_std[None, :] / W
        gb = p[K * d: K * d + K] / W
        loss = p[K * d + K] / W + 0.5 * float(np.sum(l2w[None, :] * C * C))
        G = G + l2w[None, :] * C
        return loss, (np.concatenate([G.ravel(), gb]) if fit_intercept else G.ravel())

    return f


def _parse_thresholds(th):
    """`thresholds` as a list of floats.  The widget hands every non-numeric entry over as the STRING the user
    typed (orangecontrib/spark/utils/gui_utils.py:78-94), e.g. "[0.5, 0.5]" or "0.3,0.7"."""
    if th is None:
        return None
    if isinstance(th, str):
        txt = th.strip().strip("[]()")
        if not txt or txt == "None":
            return None
        try:
            th = [float(v) for v in txt.replace(";", ",").split(",") if v.strip()]
        except ValueError:
            raise ValueError("thresholds: cannot read %r as a list of numbers" % th)
    out = [float(v) for v in th]
    if any(v < 0 for v in out) or sum(1 for v in out if v == 0) > 1:
        raise ValueError("thresholds must be >= 0 with at most one zero, got %r" % (out,))
    return out


_LR_PARAMS = (
    ("featuresCol", "features column name.", TC.toString),
    ("labelCol", "label column name.", TC.toString),
    ("predictionCol", "prediction column name.", TC.toString),
    ("maxIter", "max number of iterations (>= 0).", TC.toInt),
    ("regParam", "regularization parameter (>= 0).", TC.toFloat),
    ("elasticNetParam", "the ElasticNet mixing parameter, in range [0, 1]. For alpha = 0, the penalty is an L2 "
                        "penalty. For alpha = 1, it is an L1 penalty.", TC.toFloat),
    ("tol", "the convergence tolerance for iterative algorithms (>= 0).", TC.toFloat),
    ("fitIntercept", "whether to fit an intercept term.", TC.toBoolean),
    ("threshold", "Threshold in binary classification prediction, in range [0, 1].", TC.toFloat),
    ("thresholds", "Thresholds in multi-class classification to adjust the probability of predicting each class.",
     TC.identity),
    ("probabilityCol", "Column name for predicted class conditional probabilities.", TC.toString),
    ("rawPredictionCol", "raw prediction (a.k.a. confidence) column name.", TC.toString),
    ("standardization", "whether to standardize the training features before fitting the model.", TC.toBoolean),
    ("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
     TC.toString),
    ("aggregationDepth", "suggested depth for treeAggregate (>= 2); accepted for compatibility, the reduction is "
                         "one NCCL all-reduce.", TC.toInt),
    ("family", "The name of family which is a description of the label distribution to be used in the model. "
               "Supported options: auto, binomial, multinomial", TC.toString),
    ("lowerBoundsOnCoefficients", "not supported (bound-constrained optimisation is out of scope).", TC.identity),
    ("upperBoundsOnCoefficients", "not supported (bound-constrained optimisation is out of scope).", TC.identity),
    ("lowerBoundsOnIntercepts", "not supported (bound-constrained optimisation is out of scope).", TC.identity),
    ("upperBoundsOnIntercepts", "not supported (bound-constrained optimisation is out of scope).", TC.identity),
    ("maxBlockSizeInMB", "accepted for pyspark compatibility; ignored.", TC.toFloat),
)

_LR_DEFAULTS = dict(featuresCol="features", labelCol="label", predictionCol="prediction", maxIter=100, regParam=0.0,
                    elasticNetParam=0.0, tol=1e-6, fitIntercept=True, threshold=0.5, probabilityCol="probability",
                    rawPredictionCol="rawPrediction", standardization=True, aggregationDepth=2, family="auto",
                    maxBlockSizeInMB=0.0)


class LogisticRegressionSummary:
    def __init__(self, objective_history, total_iterations, n_evals):
        self.objectiveHistory = list(objective_history)
        self.totalIterations = total_iterations
        self.lossGradPasses = n_evals


@make_accessors
class LogisticRegressionModel(Model):
    """Model fitted by LogisticRegression."""

    _declared = tuple(p for p in _LR_PARAMS if p[0] in ("featuresCol", "labelCol", "predictionCol", "threshold",
                                                        "thresholds", "probabilityCol", "rawPredictionCol"))

    def __init__(self, coefficient_matrix=None, intercept_vector=None, num_classes=2, summary=None):
        super().__init__()
        self._setDefault(**{k: v for k, v in _LR_DEFAULTS.items() if any(k == p[0] for p in self._declared)})
        self.coefficientMatrix = None if coefficient_matrix is None else np.atleast_2d(
            np.asarray(coefficient_matrix, dtype=np.float64))
        self.interceptVector = None if intercept_vector is None else np.atleast_1d(
            np.asarray(intercept_vector, dtype=np.float64))
        self.numClasses = int(num_classes)
        self._summary = summary

    @property
    def numFeatures(self):
        return self.coefficientMatrix.shape[1]

    @property
    def coefficients(self):
        if self.coefficientMatrix.shape[0] != 1:
            raise RuntimeError("Multinomial models contain a matrix of coefficients, use coefficientMatrix instead.")
        return self.coefficientMatrix[0]

    @property
    def intercept(self):
        if self.interceptVector.shape[0] != 1:
            raise RuntimeError("Multinomial models contain a vector of intercepts, use interceptVector instead.")
        return float(self.interceptVector[0])

    @property
    def hasSummary(self):
        return self._summary is not None

    @property
    def summary(self):
        if self._summary is None:
            raise RuntimeError("No training summary available for this LogisticRegressionModel")
        return self._summary

    def _transform(self, dataset):
        torch = _torch()
        N.require_cuda()
        L = N.lib()
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("featuresCol")))
        if X.d != self.numFeatures:
            raise ValueError("model has %d features, data has %d" % (self.numFeatures, X.d))
        dev = X.data.device
        binary = self.coefficientMatrix.shape[0] == 1
        nc = 2 if binary else self.numClasses
        beta = torch.from_numpy(np.ascontiguousarray(self.coefficientMatrix)).to(dev)
        icpt = torch.from_numpy(np.ascontiguousarray(self.interceptVector)).to(dev)
        raw = torch.empty((X.n, nc), dtype=torch.float32, device=dev)
        prob = torch.empty((X.n, nc), dtype=torch.float32, device=dev)
        pred = torch.empty(X.n, dtype=torch.float32, device=dev)
        N.check(L.b2f_lr_predict(N.ptr(X.data), X.n, X.d, X.ld, nc, N.ptr(beta), N.ptr(icpt),
                                 float(self.getOrDefault("threshold")), N.ptr(raw), N.ptr(prob), N.ptr(pred),
                                 N.stream_ptr()), "b2f_lr_predict")
        th = _parse_thresholds(self.getOrDefault("thresholds"))
        if th is not None:  # argmax of p_k / t_k (A.3)
            t = torch.tensor(th, dtype=torch.float32, device=dev)
            if t.numel() != nc:
                raise ValueError("thresholds must have length %d" % nc)
            pred = torch.argmax(prob / t[None, :], dim=1).to(torch.float32)
        out = frame
        for name, val in ((self.getOrDefault("rawPredictionCol"), raw), (self.getOrDefault("probabilityCol"), prob),
                          (self.getOrDefault("predictionCol"), pred)):
            if name:
                out = out.withColumn(name, val)
        return out


@make_accessors
class LogisticRegression(Estimator):
    """Logistic regression. This class supports multinomial logistic (softmax) and binomial
    logistic regression, fitted by L-BFGS / OWL-QN over fused B200 loss+gradient passes.
    """

    _declared = _LR_PARAMS

    def __init__(self, featuresCol="features", labelCol="label", predictionCol="prediction", maxIter=100,
                 regParam=0.0, elasticNetParam=0.0, tol=1e-6, fitIntercept=True, threshold=0.5, thresholds=None,
                 probabilityCol="probability", rawPredictionCol="rawPrediction", standardization=True,
                 weightCol=None, aggregationDepth=2, family="auto", lowerBoundsOnCoefficients=None,
                 upperBoundsOnCoefficients=None, lowerBoundsOnIntercepts=None, upperBoundsOnIntercepts=None,
                 maxBlockSizeInMB=0.0):
        super().__init__()
        self._setDefault(**_LR_DEFAULTS)
        kw = dict(locals())
        kw.pop("self")
        kw.pop("__class__", None)
        self._set(**kw)

    def setParams(self, **kw):
        return self._set(**kw)

    def _fit(self, dataset):
        torch = _torch()
        g = self.getOrDefault
        for b in ("lowerBoundsOnCoefficients", "upperBoundsOnCoefficients", "lowerBoundsOnIntercepts",
                  "upperBoundsOnIntercepts"):
            if g(b) is not None:
                raise NotImplementedError("LogisticRegression: %s (bound-constrained fit) is not supported" % b)
        reg, alpha, tol, max_iter = g("regParam"), g("elasticNetParam"), g("tol"), g("maxIter")
        if reg < 0 or not (0.0 <= alpha <= 1.0) or tol < 0 or max_iter < 0:
            raise ValueError("LogisticRegression: need regParam >= 0, 0 <= elasticNetParam <= 1, tol >= 0, maxIter >= 0")
        family = g("family")
        if family not in ("auto", "binomial", "multinomial"):
            raise ValueError("LogisticRegression: unsupported family %r" % family)
        fit_icpt, standardize = g("fitIntercept"), g("standardization")
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(g("featuresCol")))
        y = to_device_vector(frame.col(g("labelCol")), X.n)
        w = to_device_vector(frame.col(g("weightCol")), X.n) if g("weightCol") else None
        comm = Comm()
        L = N.lib()
        dev = X.data.device
        d = X.d
        # ---- pass 0: summarizer (A.3)
        mean, std, W = column_stats(X, w, comm)
        ymax = (y.max() if X.n else torch.tensor(0.0, device=dev)).reshape(1).double()
        comm.all_reduce_max(ymax)
        num_classes = int(ymax.item()) + 1
        nb = max(num_classes, 2)
        hist_t = torch.zeros(nb, dtype=torch.float64, device=dev)
        bad = torch.zeros(1, dtype=torch.int32, device=dev)
        N.check(L.b2f_label_hist(N.ptr(y), N.ptr(w), X.n, nb, N.ptr(hist_t), N.ptr(bad), N.stream_ptr()),
                "b2f_label_hist")
        comm.all_reduce_sum(hist_t)
        comm.all_reduce_max(bad)  # a bad label in ONE shard must raise on every rank, not strand the others
        if int(bad.item()):
            raise ValueError("LogisticRegression: labels must be non-negative integers (as doubles)")
        hist = hist_t.cpu().numpy()
        binomial = (num_classes <= 2) if family == "auto" else (family == "binomial")
        if binomial and num_classes > 2:
            raise ValueError("Binomial family only supports 1 or 2 outcome classes but found %d." % num_classes)
        K = 1 if binomial else nb
        if not binomial and K > 32:
            raise NotImplementedError("LogisticRegression: at most 32 classes are supported, got %d" % K)
        if fit_icpt and np.count_nonzero(hist) == 1:
            # MLlib: all labels equal and fitIntercept -> no training; zero coefficients, infinite intercept
            # (+inf on the constant class; binomial: +inf when every label is 1, -inf when every label is 0)
            j = int(np.argmax(hist))
            if binomial:
                b = np.array([np.inf if num_classes == 2 else -np.inf])
            else:
                b = np.zeros(K)
                b[j] = np.inf
            summ = LogisticRegressionSummary([], 0, 0)
            summ.iterates, summ.passesAtState = [], []
            summ.featuresStd, summ.featuresMean, summ.weightSum, summ.histogram = std, mean, W, hist
            model = LogisticRegressionModel(np.zeros((K, d)), b, 2 if binomial else K, summ)
            model._set(featuresCol=g("featuresCol"), labelCol=g("labelCol"), predictionCol=g("predictionCol"),
                       threshold=g("threshold"), thresholds=g("thresholds"), probabilityCol=g("probabilityCol"),
                       rawPredictionCol=g("rawPredictionCol"))
            return model
        active = (std != 0.0).astype(np.float64)
        inv_std = np.where(std != 0.0, 1.0 / np.where(std != 0.0, std, 1.0), 0.0)
        lam2, lam1 = reg * (1.0 - alpha), reg * alpha
        l2w = np.full(d, lam2) if standardize else lam2 * inv_std ** 2
        l1w = np.full(d, lam1) if standardize else lam1 * inv_std
        eng = CudaLogRegEngine(X, y, w, 2 if binomial else K)
        f = make_objective(eng, comm, K, d, inv_std, l2w, fit_icpt, active)
        x0 = np.zeros(K * d + (K if fit_icpt else 0))
        if fit_icpt:
            if binomial:
                x0[K * d] = np.log(hist[1] / hist[0])
            else:
                raw = np.log1p(hist[:K])
                x0[K * d:] = raw - raw.mean()
        if lam1 == 0.0:
            opt = LBFGS(max_iter, 10, tol, resolution=1e-8)
        else:
            opt = OWLQN(max_iter, 10, np.concatenate([np.tile(l1w, K), np.zeros(K if fit_icpt else 0)]), tol,
                        resolution=1e-8)
        trail = opt.run(f, x0)
        last = trail[-1]
        Cs = last.x[: K * d].reshape(K, d) * active[None, :]
        b = np.array(last.x[K * d:]) if fit_icpt else np.zeros(K)
        coef = Cs * inv_std[None, :]
        if not binomial:
            if fit_icpt:
                b = b - b.mean()
            if reg == 0.0:
                coef = coef - coef.mean(axis=0, keepdims=True)
        summ = LogisticRegressionSummary([s.adj_value for s in trail], last.iter, last.n_evals)
        summ.iterates = [s.x.copy() for s in trail]
        summ.passesAtState = [s.n_evals for s in trail]
        summ.featuresStd, summ.featuresMean, summ.weightSum, summ.histogram = std, mean, W, hist
        model = LogisticRegressionModel(coef, b, 2 if binomial else K, summ)
        model._set(featuresCol=g("featuresCol"), labelCol=g("labelCol"), predictionCol=g("predictionCol"),
                   threshold=g("threshold"), thresholds=g("thresholds"), probabilityCol=g("probabilityCol"),
                   rawPredictionCol=g("rawPredictionCol"))
        return model
