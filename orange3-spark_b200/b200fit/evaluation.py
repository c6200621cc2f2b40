"""Evaluators behind the reference's "Evaluation" widget.

The widget lists every non-abstract class of its `module` that has `evaluate` (reference:
orangecontrib/spark/utils/ml_api_utils.py:59-62), reads the metric names out of the doc of the `metricName`
param -- the text between the last '(' and ')' split on '|' (orangecontrib/spark/widgets/ml/spark_ml_evaluation.py:43)
-- and evaluates once per metric with `metricName` overridden in the param map (:49-52; it calls `.apply`, a method
pyspark evaluators do not have: the widget is broken upstream.  `apply` is provided here as an alias of `evaluate`
so the widget works as written).

  ClusteringEvaluator                 silhouette (squared Euclidean, Spark's per-cluster-statistics form): two
                                      hand-written CUDA passes, b2f_silhouette_stats / b2f_silhouette_score
  BinaryClassificationEvaluator       areaUnderROC | areaUnderPR, exact curves (Spark's numBins = 0); the sort is a
                                      library call (torch.sort), the rest cumulative sums
  MulticlassClassificationEvaluator   f1 | accuracy | weightedPrecision | weightedRecall from the confusion matrix

Row-sharded multi-GPU: the silhouette statistics and the confusion matrix are all-reduced; the ranking metrics need a
global sort and are single-rank only.
"""
from abc import ABCMeta, abstractmethod

import numpy as np

from . import _native as N
from .dist import Comm
from .feature import _torch, to_device_matrix, to_device_vector
from .frame import as_frame
from .param import Params
from .param import TypeConverters as TC
from .param import make_accessors


class Evaluator(Params, metaclass=ABCMeta):
    """pyspark.ml.evaluation.Evaluator protocol: evaluate(dataset, params=None), isLargerBetter().  Abstract, as in
    pyspark: the widget's discovery skips abstract classes (orangecontrib/spark/utils/ml_api_utils.py:59-62)."""

    @abstractmethod
    def _evaluate(self, dataset):
        raise NotImplementedError

    def evaluate(self, dataset, params=None):
        if params:
            return self.copy(params)._evaluate(dataset)
        return self._evaluate(dataset)

    # the reference widget calls method_instance.apply(df, paramMap) (spark_ml_evaluation.py:52)
    apply = evaluate

    def isLargerBetter(self):
        return True


def _weights(frame, name, n):
    return to_device_vector(frame.col(name), n) if name else None


@make_accessors
class ClusteringEvaluator(Evaluator):
    """Evaluator for clustering results: the silhouette measure with squared Euclidean distance, computed from
    per-cluster statistics as Spark does (no pairwise distances)."""

    _declared = (
        ("predictionCol", "prediction column name.", TC.toString),
        ("featuresCol", "features column name.", TC.toString),
        ("metricName", "metric name in evaluation (silhouette)", TC.toString),
        ("distanceMeasure", "The distance measure. Supported options: 'squaredEuclidean'.", TC.toString),
        ("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
         TC.toString),
    )

    def __init__(self, predictionCol="prediction", featuresCol="features", metricName="silhouette",
                 distanceMeasure="squaredEuclidean", weightCol=None):
        super().__init__()
        self._setDefault(predictionCol="prediction", featuresCol="features", metricName="silhouette",
                         distanceMeasure="squaredEuclidean")
        self._set(predictionCol=predictionCol, featuresCol=featuresCol, metricName=metricName,
                  distanceMeasure=distanceMeasure, weightCol=weightCol)

    def _evaluate(self, dataset):
        N.require_cuda()
        torch = _torch()
        if self.getOrDefault("metricName") != "silhouette":
            raise ValueError("ClusteringEvaluator: unsupported metricName %r" % self.getOrDefault("metricName"))
        if self.getOrDefault("distanceMeasure") != "squaredEuclidean":
            raise ValueError("ClusteringEvaluator: only distanceMeasure='squaredEuclidean' is implemented")
        frame = as_frame(dataset)
        X = to_device_matrix(frame.col(self.getOrDefault("featuresCol")))
        lab = frame.col(self.getOrDefault("predictionCol"))
        lab = lab if isinstance(lab, torch.Tensor) else torch.from_numpy(np.asarray(lab))
        lab = lab.to(device=X.data.device, dtype=torch.int32).contiguous()
        comm = Comm()
        if comm.any_rank(lab.shape[0] != X.n):  # tested on every rank together: nobody raises alone
            raise ValueError("prediction column has %d rows, features %d (this or another rank)" % (lab.shape[0], X.n))
        wname = self.getOrDefault("weightCol") if self.isDefined("weightCol") else None
        w = _weights(frame, wname, X.n)
        kmax = torch.tensor([int(lab.max().item()) + 1 if X.n else 0], dtype=torch.int64, device=X.data.device)
        comm.all_reduce_max(kmax)
        k = int(kmax.item())
        if k < 2:
            raise ValueError("ClusteringEvaluator: the silhouette needs at least two clusters")
        L = N.lib()
        stats = torch.zeros(k * (X.d + 2), dtype=torch.float64, device=X.data.device)
        bad = torch.zeros(1, dtype=torch.int32, device=X.data.device)
        s = N.stream_ptr()
        N.check(L.b2f_silhouette_stats(N.ptr(X.data), X.n, X.d, X.ld, N.ptr(lab), N.ptr(w), k, N.ptr(stats),
                                       N.ptr(bad), s), "b2f_silhouette_stats")
        comm.all_reduce_sum(stats)
        comm.all_reduce_max(bad)
        if int(bad.item()):
            raise ValueError("ClusteringEvaluator: negative cluster index in %r" % self.getOrDefault("predictionCol"))
        scratch = torch.empty(int(L.b2f_silhouette_scratch_bytes(k, X.d)), dtype=torch.uint8, device=X.data.device)
        out = torch.zeros(2, dtype=torch.float64, device=X.data.device)
        N.check(L.b2f_silhouette_score(N.ptr(X.data), X.n, X.d, X.ld, N.ptr(lab), N.ptr(w), k, N.ptr(stats),
                                       N.ptr(scratch), N.ptr(out), s), "b2f_silhouette_score")
        comm.all_reduce_sum(out)
        tot, wsum = (float(v) for v in out.cpu())
        return tot / wsum


def _binary_areas(score, y, w):
    """Exact areas under the ROC and precision-recall curves (one point per distinct score)."""
    torch = _torch()
    score = score.double()
    order = torch.sort(score, descending=True, stable=True).indices
    s = score[order]
    ww = torch.ones_like(s) if w is None else w.double()[order]
    pos = ww * (y[order] > 0.5)
    neg = ww - pos
    last = torch.ones_like(s, dtype=torch.bool)
    last[:-1] = s[1:] != s[:-1]
    tp = torch.cumsum(pos, 0)[last]
    fp = torch.cumsum(neg, 0)[last]
    P, Nn = tp[-1], fp[-1]
    if float(P) <= 0 or float(Nn) <= 0:
        raise ValueError("BinaryClassificationEvaluator needs both classes present")
    zero, one = torch.zeros(1, dtype=s.dtype, device=s.device), torch.ones(1, dtype=s.dtype, device=s.device)
    tpr = torch.cat([zero, tp / P, one])
    fpr = torch.cat([zero, fp / Nn, one])
    roc = float(((fpr[1:] - fpr[:-1]) * (tpr[1:] + tpr[:-1]) * 0.5).sum())
    rec = torch.cat([zero, tp / P])
    prec = tp / (tp + fp)
    prec = torch.cat([prec[:1], prec])
    pr = float(((rec[1:] - rec[:-1]) * (prec[1:] + prec[:-1]) * 0.5).sum())
    return roc, pr


@make_accessors
class BinaryClassificationEvaluator(Evaluator):
    """Evaluator for binary classification, which expects input columns rawPrediction, label and an optional weight
    column.  The rawPrediction column can be of type double (binary 0/1 prediction, or probability of label 1) or of
    type vector (length-2 vector of raw predictions, scores, or label probabilities)."""

    _declared = (
        ("rawPredictionCol", "raw prediction (a.k.a. confidence) column name.", TC.toString),
        ("labelCol", "label column name.", TC.toString),
        ("metricName", "metric name in evaluation (areaUnderROC|areaUnderPR)", TC.toString),
        ("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
         TC.toString),
        ("numBins", "Number of bins to down-sample the curves (ROC curve, PR curve) in area computation. Only 0 "
                    "(exact curves, no down-sampling) is implemented; other values are accepted and ignored.",
         TC.toInt),
    )

    def __init__(self, rawPredictionCol="rawPrediction", labelCol="label", metricName="areaUnderROC", weightCol=None,
                 numBins=1000):
        super().__init__()
        self._setDefault(rawPredictionCol="rawPrediction", labelCol="label", metricName="areaUnderROC", numBins=1000)
        self._set(rawPredictionCol=rawPredictionCol, labelCol=labelCol, metricName=metricName, weightCol=weightCol,
                  numBins=numBins)

    def _evaluate(self, dataset):
        N.require_cuda()
        torch = _torch()
        metric = self.getOrDefault("metricName")
        if metric not in ("areaUnderROC", "areaUnderPR"):
            raise ValueError("BinaryClassificationEvaluator: unsupported metricName %r" % metric)
        if Comm().world > 1:
            raise NotImplementedError("BinaryClassificationEvaluator needs a global sort: single rank only")
        frame = as_frame(dataset)
        raw = frame.col(self.getOrDefault("rawPredictionCol"))
        raw = raw.data[:, : raw.d] if hasattr(raw, "ld") else raw
        raw = raw if isinstance(raw, torch.Tensor) else torch.from_numpy(np.asarray(raw))
        raw = raw.cuda()
        score = raw[:, 1] if raw.dim() == 2 and raw.shape[1] >= 2 else raw.reshape(-1)
        n = score.shape[0]
        y = to_device_vector(frame.col(self.getOrDefault("labelCol")), n)
        wname = self.getOrDefault("weightCol") if self.isDefined("weightCol") else None
        roc, pr = _binary_areas(score, y, _weights(frame, wname, n))
        return roc if metric == "areaUnderROC" else pr


@make_accessors
class MulticlassClassificationEvaluator(Evaluator):
    """Evaluator for multiclass classification, which expects input columns prediction, label and an optional weight
    column."""

    _declared = (
        ("predictionCol", "prediction column name.", TC.toString),
        ("labelCol", "label column name.", TC.toString),
        ("metricName", "metric name in evaluation (f1|accuracy|weightedPrecision|weightedRecall)", TC.toString),
        ("weightCol", "weight column name. If this is not set or empty, we treat all instance weights as 1.0.",
         TC.toString),
    )

    def __init__(self, predictionCol="prediction", labelCol="label", metricName="f1", weightCol=None):
        super().__init__()
        self._setDefault(predictionCol="prediction", labelCol="label", metricName="f1")
        self._set(predictionCol=predictionCol, labelCol=labelCol, metricName=metricName, weightCol=weightCol)

    def _evaluate(self, dataset):
        N.require_cuda()
        torch = _torch()
        metric = self.getOrDefault("metricName")
        if metric not in ("f1", "accuracy", "weightedPrecision", "weightedRecall"):
            raise ValueError("MulticlassClassificationEvaluator: unsupported metricName %r" % metric)
        frame = as_frame(dataset)
        pred = frame.col(self.getOrDefault("predictionCol"))
        pred = (pred if isinstance(pred, torch.Tensor) else torch.from_numpy(np.asarray(pred))).cuda().long()
        n = pred.shape[0]
        y = to_device_vector(frame.col(self.getOrDefault("labelCol")), n).long()
        wname = self.getOrDefault("weightCol") if self.isDefined("weightCol") else None
        w = _weights(frame, wname, n)
        comm = Comm()
        kk = torch.stack([pred.max(), y.max()]).max().reshape(1) + 1 if n else torch.zeros(1, dtype=torch.long).cuda()
        comm.all_reduce_max(kk)
        K = int(kk.item())
        cm = torch.bincount(y * K + pred, weights=None if w is None else w.double(), minlength=K * K).double()
        comm.all_reduce_sum(cm)
        cm = cm.reshape(K, K).cpu().numpy()
        tot, tp = cm.sum(), np.diag(cm)
        by_label, by_pred = cm.sum(axis=1), cm.sum(axis=0)
        prec = np.where(by_pred > 0, tp / np.where(by_pred > 0, by_pred, 1.0), 0.0)
        rec = np.where(by_label > 0, tp / np.where(by_label > 0, by_label, 1.0), 0.0)
        f1 = np.where(prec + rec > 0, 2 * prec * rec / np.where(prec + rec > 0, prec + rec, 1.0), 0.0)
        wt = by_label / tot
        return {"accuracy": float(tp.sum() / tot), "weightedPrecision": float((wt * prec).sum()),
                "weightedRecall": float((wt * rec).sum()), "f1": float((wt * f1).sum())}[metric]
